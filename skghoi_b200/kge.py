"""TransH / MarginLoss / NegativeSampling with the reference's interfaces on the CUDA library.

Mirrors heads/TransH/TransH.py:10-132, heads/MarginLoss.py:12-41, heads/NegativeSampling.py:6-63
and the OpenKE base classes they inherit (OpenKE/openke/module/BaseModule.py:7-54,
model/Model.py, strategy/Strategy.py, loss/Loss.py): same class names, constructor signatures,
attribute / state_dict names (``zero_const``, ``pi_const``, ``ent_embeddings`` ...), so checkpoints
and ``isinstance`` checks behave.  The arithmetic is csrc/transh.cu; CUDA tensors only.
"""
import json
import os

import torch
import torch.nn as nn

from . import _lib


class BaseModule(nn.Module):
    """OpenKE/openke/module/BaseModule.py:7-54."""

    def __init__(self):
        super().__init__()
        self.zero_const = nn.Parameter(torch.Tensor([0]), requires_grad=False)
        self.pi_const = nn.Parameter(torch.Tensor([3.14159265358979323846]), requires_grad=False)

    def load_checkpoint(self, path):
        self.load_state_dict(torch.load(os.path.join(path)))
        self.eval()

    def save_checkpoint(self, path):
        torch.save(self.state_dict(), path)

    def load_parameters(self, path):
        with open(path, "r") as f:
            parameters = json.loads(f.read())
        self.set_parameters(parameters)

    def save_parameters(self, path):
        with open(path, "w") as f:
            f.write(json.dumps(self.get_parameters("list")))

    def get_parameters(self, mode="numpy", param_dict=None):
        all_param_dict = self.state_dict()
        if param_dict is None:
            param_dict = all_param_dict.keys()
        res = {}
        for param in param_dict:
            if mode == "numpy":
                res[param] = all_param_dict[param].cpu().numpy()
            elif mode == "list":
                res[param] = all_param_dict[param].cpu().numpy().tolist()
            else:
                res[param] = all_param_dict[param]
        return res

    def set_parameters(self, parameters):
        parameters = {k: torch.Tensor(v) for k, v in parameters.items()}
        self.load_state_dict(parameters, strict=False)
        self.eval()


class Model(BaseModule):
    """OpenKE/openke/module/model/Model.py."""

    def __init__(self, ent_tot, rel_tot):
        super().__init__()
        self.ent_tot = ent_tot
        self.rel_tot = rel_tot

    def forward(self):
        raise NotImplementedError

    def predict(self):
        raise NotImplementedError


class Strategy(BaseModule):
    """OpenKE/openke/module/strategy/Strategy.py."""


class Loss(BaseModule):
    """OpenKE/openke/module/loss/Loss.py."""


def _idx(t, device):
    return t.to(device=device, dtype=torch.int64).contiguous()


class _TransHScore(torch.autograd.Function):
    """(ent, rel, nrm, head, relation, tail) -> (h_, r, r_norm, t_, score)."""

    @staticmethod
    def forward(ctx, ent, rel, nrm, head, relation, tail, p_norm, norm_flag, want_rows):
        _lib.require_cuda(ent, rel, nrm, head, relation, tail)
        dev = ent.device
        T = head.numel()
        dim = ent.shape[1]
        e, r, w = ent.contiguous().float(), rel.contiguous().float(), nrm.contiguous().float()
        score = torch.empty(T, dtype=torch.float32, device=dev)
        if want_rows:
            h_o, r_o, w_o, t_o = (torch.empty(T, dim, dtype=torch.float32, device=dev) for _ in range(4))
        else:
            h_o = r_o = w_o = t_o = None
        with torch.cuda.device(dev):
            _lib.call("skg_transh_forward", _lib.ptr(e), _lib.ptr(r), _lib.ptr(w), e.shape[0], r.shape[0], dim,
                      _lib.ptr(head), _lib.ptr(relation), _lib.ptr(tail), T, int(p_norm), int(bool(norm_flag)),
                      _lib.ptr(h_o), _lib.ptr(r_o), _lib.ptr(w_o), _lib.ptr(t_o), _lib.ptr(score),
                      _lib.stream_ptr(dev))
        ctx.save_for_backward(e, r, w, head, relation, tail)
        ctx.cfg = (int(p_norm), int(bool(norm_flag)))
        if not want_rows:
            h_o, r_o, w_o, t_o = (torch.empty(0, device=dev) for _ in range(4))
            ctx.mark_non_differentiable(h_o, r_o, w_o, t_o)
        return h_o, r_o, w_o, t_o, score

    @staticmethod
    def backward(ctx, g_h, g_r, g_w, g_t, g_score):
        e, r, w, head, relation, tail = ctx.saved_tensors
        p_norm, norm_flag = ctx.cfg
        dev = e.device
        d_e, d_r, d_w = torch.zeros_like(e), torch.zeros_like(r), torch.zeros_like(w)
        if g_score is not None:
            gs = g_score.contiguous().float()
            with torch.cuda.device(dev):
                _lib.call("skg_transh_backward", _lib.ptr(e), _lib.ptr(r), _lib.ptr(w), e.shape[0], r.shape[0],
                          e.shape[1], _lib.ptr(head), _lib.ptr(relation), _lib.ptr(tail), head.numel(), p_norm,
                          norm_flag, _lib.ptr(gs), _lib.ptr(d_e), _lib.ptr(d_r), _lib.ptr(d_w), _lib.stream_ptr(dev))
        # gradients flowing into the raw gathered rows (only the regulariser reads them)
        for g, table, index in ((g_h, d_e, head), (g_t, d_e, tail), (g_r, d_r, relation), (g_w, d_w, relation)):
            if g is not None and g.numel() > 0:
                table.index_add_(0, index, g.float())
        return d_e, d_r, d_w, None, None, None, None, None, None


class TransH(Model):
    """heads/TransH/TransH.py:8-132 (constructor TransH.py:10-54, forward TransH.py:87-106)."""

    def __init__(self, ent_tot, rel_tot, dim=50, p_norm=1, norm_flag=True, margin=None, epsilon=None):
        super().__init__(ent_tot, rel_tot)
        self.dim = dim
        self.margin = margin
        self.epsilon = epsilon
        self.norm_flag = norm_flag
        self.p_norm = p_norm
        self.ent_embeddings = nn.Embedding(self.ent_tot, self.dim)
        self.rel_embeddings = nn.Embedding(self.rel_tot, self.dim)
        self.norm_vector = nn.Embedding(self.rel_tot, self.dim)
        if margin is None or epsilon is None:
            nn.init.xavier_uniform_(self.ent_embeddings.weight.data)
            nn.init.xavier_uniform_(self.rel_embeddings.weight.data)
            nn.init.xavier_uniform_(self.norm_vector.weight.data)
        else:
            self.embedding_range = nn.Parameter(torch.Tensor([(self.margin + self.epsilon) / self.dim]),
                                                requires_grad=False)
            for emb in (self.ent_embeddings, self.rel_embeddings, self.norm_vector):
                nn.init.uniform_(tensor=emb.weight.data, a=-self.embedding_range.item(),
                                 b=self.embedding_range.item())
        if margin is not None:
            self.margin = nn.Parameter(torch.Tensor([margin]), requires_grad=False)
            self.margin_flag = True
        else:
            self.margin_flag = False

    def _score(self, head, relation, tail, want_rows):
        dev = self.ent_embeddings.weight.device
        head, relation, tail = _idx(head, dev), _idx(relation, dev), _idx(tail, dev)
        if not (head.shape == relation.shape == tail.shape):
            raise RuntimeError("head, relation and tail must have the same shape ('normal' mode, TransH.py:88)")
        return _TransHScore.apply(self.ent_embeddings.weight, self.rel_embeddings.weight, self.norm_vector.weight,
                                  head.reshape(-1), relation.reshape(-1), tail.reshape(-1), self.p_norm,
                                  self.norm_flag, want_rows)

    def forward(self, head, relation, tail):
        if self.margin_flag:
            return self.margin - self._score(head, relation, tail, False)[4]     # TransH.py:103-104
        return self._score(head, relation, tail, True)                           # TransH.py:106

    def regularization(self, head, relation, tail):
        """TransH.py:108-120."""
        h = self.ent_embeddings(head); t = self.ent_embeddings(tail)
        r = self.rel_embeddings(relation); r_norm = self.norm_vector(relation)
        return (torch.mean(h ** 2) + torch.mean(t ** 2) + torch.mean(r ** 2) + torch.mean(r_norm ** 2)) / 4

    def l3_regularization(self, head, relation, tail):
        return self.ent_embeddings.weight.norm(p=3) ** 3 + self.rel_embeddings.weight.norm(p=3) ** 3

    def predict(self, head, relation, tail):
        """The reference's ``predict`` (TransH.py:125-131) references an undefined name; the evident
        intent -- entities plus score, sign-flipped under margin mode -- is implemented."""
        h_, _, _, t_, score = self._score(head, relation, tail, True)
        if self.margin_flag:
            score = self.margin - score
        return h_, t_, score


class _MarginLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p, n, margin, adv_t, use_adv):
        _lib.require_cuda(p, n)
        rows, cols = n.shape
        pc, nc = p.contiguous().float(), n.contiguous().float()
        loss = torch.empty(1, dtype=torch.float32, device=p.device)
        need = p.requires_grad or n.requires_grad
        d_p = torch.empty_like(pc) if need else None
        d_n = torch.empty_like(nc) if need else None
        with torch.cuda.device(p.device):
            _lib.call("skg_margin_loss", _lib.ptr(pc), _lib.ptr(nc), rows, cols, float(margin), float(adv_t),
                      int(use_adv), _lib.ptr(loss), _lib.ptr(d_p), _lib.ptr(d_n), 1.0, _lib.stream_ptr(p.device))
        if need:
            ctx.save_for_backward(d_p, d_n)
        return loss

    @staticmethod
    def backward(ctx, g):
        d_p, d_n = ctx.saved_tensors
        return d_p * g, d_n * g, None, None, None


class MarginLoss(Loss):
    """heads/MarginLoss.py:12-41."""

    def __init__(self, adv_temperature=None, margin=6.0):
        super().__init__()
        self.margin = nn.Parameter(torch.Tensor([margin]), requires_grad=False)
        if adv_temperature is not None:
            self.adv_temperature = nn.Parameter(torch.Tensor([adv_temperature]), requires_grad=False)
            self.adv_flag = True
        else:
            self.adv_flag = False

    def get_weights(self, n_score):
        return torch.softmax(-n_score * self.adv_temperature, dim=-1).detach()

    def forward(self, p_score, n_score):
        """``mean(max(p - n, -margin)) + margin`` (MarginLoss.py:31-35) or the self-adversarial form
        (MarginLoss.py:29-30).  p_score [rows] or [rows, 1] broadcasts against n_score [rows, cols]."""
        if n_score.dim() == 1:
            n_score = n_score[:, None]
        p = p_score.reshape(-1)
        if p.numel() != n_score.shape[0] or n_score.dim() != 2:
            if p_score.shape == n_score.shape:        # element-wise case
                if self.adv_flag and n_score.shape[1] > 1:
                    # per-element positives against [rows, cols] negatives under the self-adversarial weighting: the softmax
                    # runs over each ROW's columns (MarginLoss.py:29-30), which the [rows] x [rows, cols] kernel layout cannot
                    # express -- rare shape, plain tensor expression
                    m = self.margin.to(p_score.device)
                    return (self.get_weights(n_score) * torch.max(p_score - n_score, -m)).sum(dim=-1).mean() + m
                p = p_score.reshape(-1); n_score = n_score.reshape(-1, 1)
            else:
                raise RuntimeError("p_score %s does not broadcast against n_score %s"
                                   % (tuple(p_score.shape), tuple(n_score.shape)))
        adv_t = self._scalar("adv_temperature") if self.adv_flag else 0.0
        return _MarginLossFn.apply(p, n_score, self._scalar("margin"), adv_t, self.adv_flag)

    def _scalar(self, name: str) -> float:
        """Host value of a one-element hyper-parameter tensor.  Reading it off the device is a synchronisation per loss
        call (and not allowed inside a CUDA-graph capture): read once, re-read only when the tensor was replaced or
        written in place (load_state_dict, .to(...))."""
        t = getattr(self, name)
        key = (t.data_ptr(), t._version)
        cache = self.__dict__.setdefault("_scalar_cache", {})
        hit = cache.get(name)
        if hit is None or hit[0] != key:
            hit = cache[name] = (key, float(t))
        return hit[1]

    def predict(self, p_score, n_score):
        return self.forward(p_score, n_score).cpu().data.numpy()


class NegativeSampling(Strategy):
    """heads/NegativeSampling.py:6-63."""

    def __init__(self, model=None, loss=None, batch_size=256, regul_rate=0.0, l3_regul_rate=0.0):
        super().__init__()
        self.model = model
        self.loss = loss
        self.batch_size = batch_size
        self.regul_rate = regul_rate
        self.l3_regul_rate = l3_regul_rate

    def _get_positive_score(self, score):
        half = len(score) // 2
        return score[:half].view(-1, half).permute(1, 0)                  # NegativeSampling.py:30-35

    def _get_negative_score(self, score):
        half = len(score) // 2
        return score[half:].view(-1, half).permute(1, 0)                  # NegativeSampling.py:37-40

    def regularization(self, head, relation, relation_norm, tail):
        head = torch.cat(head); tail = torch.cat(tail)
        relation = torch.cat(relation); relation_norm = torch.cat(relation_norm)
        return (torch.mean(head ** 2) + torch.mean(tail ** 2) + torch.mean(relation ** 2)
                + torch.mean(relation_norm ** 2)) / 4                     # NegativeSampling.py:42-51

    def forward(self, score, negative_scores=None, head=None, relation=None, relation_norm=None, tail=None):
        """One-argument form = the class's own contract (NegativeSampling.py:52-63): a flat score
        vector, first half positives, second half negatives.  The six-argument form is what the
        head's call site passes (head.py:233: lists of per-image positive / negative scores and the
        gathered embeddings); the reference raises TypeError there (SURVEY defect 2), here it
        evaluates the intended ``loss(cat(pos), cat(neg))``."""
        if negative_scores is not None:
            pos = torch.cat(list(score)) if isinstance(score, (list, tuple)) else score
            neg = torch.cat(list(negative_scores)) if isinstance(negative_scores, (list, tuple)) else negative_scores
            score = torch.cat([pos, neg])
        p_score = self._get_positive_score(score)
        n_score = self._get_negative_score(score)
        loss_res = self.loss(p_score, n_score)
        if self.regul_rate != 0:
            if head is None:
                raise RuntimeError("regul_rate != 0 needs the gathered embeddings (six-argument form)")
            loss_res = loss_res + self.regul_rate * self.regularization(head, relation, relation_norm, tail)
        return loss_res
