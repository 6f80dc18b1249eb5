"""Pins the stock-PyTorch RAFT harness to the unmodified reference model (CPU), and the oracle's CorrBlock
restatement inside the full 12-iteration loop.  The reference tree is taken from baseline/_ref (the copy
tools/install_ref.py makes, which travels to the GPU box) or from /root/reference; skipped when neither exists."""
import argparse
import os
import sys
import warnings

import pytest
import torch

from harness import ref_model

warnings.filterwarnings("ignore")
REF = ref_model.ref_core_dir() or "/root/reference/core"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")
@pytest.mark.parametrize("small", [False, True])
def test_harness_matches_reference_model(small):
    from harness.raft_stock import StockRAFT
    from oracle import torch_restatement as tr

    saved = list(sys.path)
    saved_mods = {k: sys.modules.pop(k) for k in ("corr", "raft", "utils", "utils.utils", "update", "extractor") if k in sys.modules}
    sys.path.insert(0, REF)
    try:
        from raft import RAFT
        torch.manual_seed(1234)
        ref = RAFT(argparse.Namespace(small=small, mixed_precision=False, alternate_corr=False)).eval()
        mine = StockRAFT(small=small).eval()
        missing = mine.load_state_dict(ref.state_dict(), strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys
        gen = torch.Generator().manual_seed(5)
        im1 = torch.rand(1, 3, 128, 160, generator=gen) * 255
        im2 = torch.rand(1, 3, 128, 160, generator=gen) * 255
        with torch.no_grad():
            lo_ref, up_ref = ref(im1, im2, iters=4, test_mode=True)
            lo, up = mine(im1, im2, lambda a, b, r: tr.TorchCorrBlock(a, b, 4, r), iters=4)
        assert up.shape == up_ref.shape == (1, 2, 128, 160)
        assert (lo - lo_ref).abs().max().item() <= 1e-4
        epe = (up - up_ref).pow(2).sum(1).sqrt().mean().item()
        assert epe <= 1e-4, epe
    finally:
        sys.path[:] = saved
        for k in ("corr", "raft", "utils", "utils.utils", "update", "extractor"):
            sys.modules.pop(k, None)
        sys.modules.update(saved_mods)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")
@pytest.mark.parametrize("alternate", [False, True])
def test_unmodified_reference_model_picks_up_the_drop_in(alternate):
    """with this repository's package directory ahead of core/ on sys.path, the reference's
    own raft.py (unmodified) binds this library's CorrBlock / AlternateCorrBlock.  There is no GPU in the build
    container, so the proof is that the forward reaches the drop-in and stops at its no-CPU-fallback check."""
    import corr as dropin
    names = ("corr", "raft", "utils", "utils.utils", "update", "extractor")
    saved = list(sys.path)
    saved_mods = {k: sys.modules.pop(k) for k in names if k in sys.modules}
    sys.modules["corr"] = dropin               # what `sys.path.insert(0, "pytorch-raft_b200")` resolves to
    sys.path.insert(1, REF)
    try:
        import raft
        assert raft.CorrBlock is dropin.CorrBlock and raft.AlternateCorrBlock is dropin.AlternateCorrBlock
        model = raft.RAFT(argparse.Namespace(small=False, mixed_precision=False, alternate_corr=alternate)).eval()
        im = torch.rand(1, 3, 128, 160) * 255
        with torch.no_grad(), pytest.raises(RuntimeError, match="CUDA tensor"):
            model(im, im, iters=1, test_mode=True)
    finally:
        sys.path[:] = saved
        for k in names:
            sys.modules.pop(k, None)
        sys.modules.update(saved_mods)
