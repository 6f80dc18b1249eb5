"""north_star's API clause, exercised for real: the reference's OWN ``core/raft.py`` — unmodified, imported from
``baseline/_ref/core`` (tools/install_ref.py) — runs on cuda:0 with this repository's ``pytorch-raft_b200/`` ahead of
``core/`` on ``sys.path`` (raft.py:8 ``from corr import CorrBlock, AlternateCorrBlock``; raft.py:104-107, 122-124),
and its 12-iteration flow is compared with the same model, same weights, same frames, running the reference's own
``corr.py`` on the same GPU.  Bar: mean end-point error <= 0.01 px (BASELINE.json north_star), at 440x1024.

The reference's AlternateCorrBlock needs its ``alt_cuda_corr`` extension, which is not built anywhere (and must
not be recompiled as our implementation); its expected flow is the reference CorrBlock's — the same mathematics
(SURVEY.md §8 a-7).
"""
import warnings

import pytest
import torch

from harness import ref_model

warnings.filterwarnings("ignore")
pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_model.available(), reason="baseline/_ref not installed (tools/install_ref.py)")]

ITERS = 12


def _frames(B, H, W, seed):
    gen = torch.Generator().manual_seed(seed)
    return ((torch.rand(B, 3, H, W, generator=gen) * 255).cuda(), (torch.rand(B, 3, H, W, generator=gen) * 255).cuda())


def _epe(a, b):
    return (a - b).pow(2).sum(1).sqrt().mean().item()


@pytest.fixture(scope="module")
def raft_modules():
    return ref_model.load_raft_module(dropin=True), ref_model.load_raft_module(dropin=False)


def test_imports_bind_the_right_corr(raft_modules):
    ours, theirs = raft_modules
    import corr as dropin
    assert ours.CorrBlock is dropin.CorrBlock and ours.AlternateCorrBlock is dropin.AlternateCorrBlock
    assert theirs.CorrBlock is not dropin.CorrBlock
    assert theirs.CorrBlock.__module__ == "corr"
    assert "baseline/_ref" in theirs.__file__ or "/root/reference" in theirs.__file__


@pytest.mark.parametrize("small", [False, True], ids=["basic", "small"])
@pytest.mark.parametrize("mixed", [False, True], ids=["fp32nets", "autocast"])
def test_unmodified_raft_flow_matches_reference_corr(raft_modules, small, mixed, monkeypatch):
    ours_mod, ref_mod = raft_modules
    torch.manual_seed(1234)
    ref = ref_mod.RAFT(ref_model.make_args(small=small, mixed_precision=mixed)).eval().cuda()
    im1, im2 = _frames(1, 440, 1024, seed=21)
    with torch.no_grad():
        lo_ref, up_ref = ref(im1, im2, iters=ITERS, test_mode=True)
    assert up_ref.shape == (1, 2, 440, 1024)
    mag = up_ref.pow(2).sum(1).sqrt().mean().item()
    assert mag > 0.05, mag      # random-init RAFT does move pixels: the comparison is not vacuous
    for alternate in (False, True):
        for precision in ("bf16", "fp32"):
            if alternate and precision == "fp32" and small:
                continue   # covered by the basic model; keeps the test short
            monkeypatch.setenv("RAFTCORR_PRECISION", precision)
            model = ours_mod.RAFT(ref_model.make_args(small=small, mixed_precision=mixed,
                                                       alternate_corr=alternate)).eval().cuda()
            model.load_state_dict(ref.state_dict(), strict=True)
            with torch.no_grad():
                lo, up = model(im1, im2, iters=ITERS, test_mode=True)
            assert up.shape == up_ref.shape and lo.shape == lo_ref.shape
            assert torch.isfinite(up).all()
            epe = _epe(up, up_ref)
            assert epe <= 0.01, dict(small=small, mixed=mixed, alternate=alternate, precision=precision, epe=epe,
                                     mean_flow=mag)


def test_unmodified_raft_under_dataparallel_and_training_mode(raft_modules, monkeypatch):
    """train.py:138 wraps the model in nn.DataParallel and calls it in training mode (list of 12 predictions,
    gradients through the correlation block): the drop-in must serve that call too."""
    ours_mod, ref_mod = raft_modules
    monkeypatch.setenv("RAFTCORR_PRECISION", "fp32")
    torch.manual_seed(7)
    ref = ref_mod.RAFT(ref_model.make_args()).cuda()
    model = ours_mod.RAFT(ref_model.make_args()).cuda()
    model.load_state_dict(ref.state_dict(), strict=True)
    ref.train(), model.train()
    ref.freeze_bn(), model.freeze_bn()
    wrapped = torch.nn.DataParallel(model, device_ids=[0])
    im1, im2 = _frames(2, 128, 192, seed=3)
    preds = wrapped(im1, im2, iters=4)
    preds_ref = ref(im1, im2, iters=4)
    assert len(preds) == len(preds_ref) == 4
    assert _epe(preds[-1], preds_ref[-1]) <= 0.01
    loss = sum(p.abs().mean() for p in preds)
    loss_ref = sum(p.abs().mean() for p in preds_ref)
    loss.backward()
    loss_ref.backward()
    g = dict(model.named_parameters())
    checked = 0
    for name, p_ref in ref.named_parameters():
        if p_ref.grad is None:
            continue
        a, b = g[name].grad, p_ref.grad
        assert a is not None, name
        scale = b.abs().max().item()
        if scale == 0:
            continue
        # recurrent instance-norm networks amplify rounding: compare at 2 % of each parameter's largest gradient
        assert (a - b).abs().max().item() <= 2e-2 * scale + 1e-6, (name, (a - b).abs().max().item(), scale)
        checked += 1
    assert checked > 50
