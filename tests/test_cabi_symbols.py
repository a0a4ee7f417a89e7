"""The C-ABI library builds, loads, and exports every symbol include/*.h declares (no compute)."""
import ctypes
import glob
import os
import re

from helpers import ROOT


def _declared():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names.update(re.findall(r"\b(skg_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_declares_entry_points():
    names = _declared()
    for must in ("skg_pairs_spatial_encode", "skg_spatial_encode_rows", "skg_transh_forward",
                 "skg_transh_backward", "skg_margin_loss", "skg_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from skghoi_b200 import build, _lib
    path = build.build()
    assert os.path.isfile(path)
    lib = ctypes.CDLL(path)
    for name in _declared():
        assert hasattr(lib, name), "missing export: " + name
    # the ctypes signature table covers the header, and vice versa
    assert set(_lib.declared_symbols()) == set(_declared())
    assert _lib.lib().skg_abi_version() == 1


def test_ops_refuse_cpu_tensors():
    """No CPU fallback: the product operators raise on non-CUDA tensors."""
    import pytest
    import torch
    from skghoi_b200 import ops, _lib
    b = torch.tensor([[0., 0., 10., 10.]])
    with pytest.raises(_lib.SkgError):
        ops.compute_spatial_ratio_encodings([b], [b], [(100, 100)])
    from skghoi_b200.kge import TransH
    m = TransH(8, 4, dim=6, p_norm=2)
    with pytest.raises(_lib.SkgError):
        m(torch.tensor([0]), torch.tensor([1]), torch.tensor([2]))
