"""Host-side logic that needs no GPU: the batched negative selection of the TransH loss (device-generator form of
head.py:936-943) on CPU tensors."""
import torch

from skghoi_b200.engine import select_negatives_device


def test_select_negatives_per_image_counts_and_validity():
    torch.manual_seed(0)
    P, nc, B = 40, 117, 4                                 # image 3 has no pairs at all
    pair_img = torch.tensor([0] * 10 + [1] * 25 + [2] * 5)
    labels = (torch.rand(P, nc) > 0.97).float()
    labels[35:] = 0; labels[36, 5] = 1
    pos, neg = select_negatives_device(labels, pair_img, B)
    flat = labels.view(-1)
    assert torch.equal(pos, torch.nonzero(flat).squeeze(1))
    assert bool((flat[neg] == 0).all()) and neg.unique().numel() == neg.numel()
    for b in range(B):
        n_pos = int((pair_img[pos // nc] == b).sum()); n_neg = int((pair_img[neg // nc] == b).sum())
        assert n_pos == n_neg, (b, n_pos, n_neg)
    # a different draw gives a different subset, same counts
    _, neg2 = select_negatives_device(labels, pair_img, B)
    assert neg2.numel() == neg.numel() and not torch.equal(neg2.sort().values, neg.sort().values)


def test_select_negatives_no_positives():
    labels = torch.zeros(6, 117)
    pos, neg = select_negatives_device(labels, torch.zeros(6, dtype=torch.int64), 1)
    assert pos.numel() == 0 and neg.numel() == 0
