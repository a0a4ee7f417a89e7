"""Shared helpers for the test-suite (parameter rebuild, tolerances)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from skghoi_b200 import synthetic  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
_cache = {}


def load_golden(name):
    if name not in _cache:
        _cache[name] = torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    return _cache[name]


def build_head(seed=42, **kw):
    """Product InteractionHead on the CPU with reference-identical initial weights."""
    from skghoi_b200.head import build_interaction_head
    return build_interaction_head(synthetic.hico_object_to_verb(), seed=seed, **kw)


def head_params(seed=42):
    key = ("params", seed)
    if key not in _cache:
        _cache[key] = {k: v.detach().clone() for k, v in build_head(seed).state_dict().items()}
    return _cache[key]


def logits_close(got, ref, tol=1e-2):
    """north_star: interaction logits within 1e-2 relative (SURVEY 8(d) operational form):
    max|d| <= tol*max|ref| and per element |d| <= tol*|ref| + tol*rms(ref)."""
    got = got.double().cpu(); ref = ref.double().cpu()
    d = (got - ref).abs()
    rms = ref.pow(2).mean().sqrt()
    ok_max = bool(d.max() <= tol * ref.abs().max())
    ok_el = bool((d <= tol * ref.abs() + tol * rms).all())
    return ok_max and ok_el, dict(max_abs=float(d.max()), ref_max=float(ref.abs().max()), rms=float(rms),
                                  worst_el=float((d - tol * ref.abs() - tol * rms).max()))
