"""Native multi-scale RoIAlign (skg_multiscale_roi_align) against the operator the reference calls: torchvision's
MultiScaleRoIAlign(['0','1','2','3'], 7, 2) (models.py:158-162, head.py:387), run on the CPU as the oracle."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _pyramid(B, C, H, W, gen):
    sizes = [((H + s - 1) // s, (W + s - 1) // s) for s in (4, 8, 16, 32)]
    feats = {str(i): torch.randn(B, C, h, w, generator=gen) for i, (h, w) in enumerate(sizes)}
    feats["pool"] = torch.randn(B, C, 2, 2, generator=gen)        # an extra map the pooler must ignore
    return feats


def _boxes(n, H, W, gen, lo, hi):
    wh = lo + (hi - lo) * torch.rand(n, 2, generator=gen)
    xy = torch.rand(n, 2, generator=gen) * torch.tensor([W, H]) - 0.25 * wh     # some boxes start left of / above the image
    return torch.cat([xy, xy + wh], dim=1).float()


@pytest.mark.parametrize("C,B", [(32, 3), (256, 2)])
def test_multiscale_roi_align_vs_torchvision(C, B):
    from torchvision.ops import MultiScaleRoIAlign as TV
    from skghoi_b200.roi import MultiScaleRoIAlign
    gen = torch.Generator().manual_seed(5 + C)
    H, W = 400, 608
    feats = _pyramid(B, C, H, W, gen)
    shapes = [(H, W)] + [(H - 40, W - 64)] * (B - 1)
    boxes = [torch.cat([_boxes(9, H, W, gen, 8, 60), _boxes(9, H, W, gen, 60, 250), _boxes(6, H, W, gen, 250, 600)])
             for _ in range(B)]
    boxes[1] = torch.cat([boxes[1], torch.tensor([[10., 10., 10., 10.], [30., 40., 30.5, 300.]])])   # zero-area, sliver
    boxes[-1] = boxes[-1][:5]
    ref = TV(["0", "1", "2", "3"], 7, 2)(feats, boxes, shapes)
    mine = MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)
    got = mine({k: v.to(DEV) for k, v in feats.items()}, [b.to(DEV) for b in boxes], shapes)
    assert got.shape == ref.shape and got.dtype == torch.float32
    torch.testing.assert_close(got.cpu(), ref, rtol=1e-5, atol=1e-5)
    # every pyramid level was exercised
    lv = TV(["0", "1", "2", "3"], 7, 2)
    lv(feats, boxes, shapes)
    assert set(lv.map_levels(boxes).tolist()) == {0, 1, 2, 3}
    # bf16 operand rows = the fp32 result rounded, written with a padded leading dimension
    rows = torch.zeros(got.shape[0], C * 49 + 8, dtype=torch.bfloat16, device=DEV)
    assert mine.pool({k: v.to(DEV) for k, v in feats.items()}, [b.to(DEV) for b in boxes], shapes, want_f32=False,
                     bf16_out=rows) is None
    assert torch.equal(rows[:, :C * 49].cpu(), got.flatten(1).to(torch.bfloat16).cpu())
    assert float(rows[:, C * 49:].abs().max()) == 0.0


def test_single_level_and_empty():
    from torchvision.ops import MultiScaleRoIAlign as TV
    from skghoi_b200.roi import MultiScaleRoIAlign
    gen = torch.Generator().manual_seed(3)
    feats = {"3": torch.randn(2, 16, 25, 38, generator=gen)}
    shapes = [(200, 304), (180, 300)]
    boxes = [_boxes(7, 200, 304, gen, 10, 150), torch.zeros(0, 4)]
    ref = TV(["3"], 7, 2)(feats, boxes, shapes)
    got = MultiScaleRoIAlign(["3"], 7, 2)({"3": feats["3"].to(DEV)}, [b.to(DEV) for b in boxes], shapes)
    torch.testing.assert_close(got.cpu(), ref, rtol=1e-5, atol=1e-5)
    none = MultiScaleRoIAlign(["3"], 7, 2)({"3": feats["3"].to(DEV)}, [torch.zeros(0, 4, device=DEV)] * 2, shapes)
    assert none.shape == (0, 16, 7, 7)


def test_roi_align_rejects_cpu_tensors():
    from skghoi_b200 import _lib
    from skghoi_b200.roi import MultiScaleRoIAlign
    with pytest.raises(_lib.SkgError):
        MultiScaleRoIAlign(["0"], 7, 2)({"0": torch.zeros(1, 4, 8, 8)}, [torch.zeros(1, 4)], [(32, 32)])
