"""raftcorr_b200.graphed: a whole 12-iteration forward of the UNMODIFIED reference model (with the drop-in) replayed
as one CUDA graph gives bit for bit what the eager forward gives, for new frames as well."""
import warnings

import pytest
import torch

import raftcorr_b200
from harness import ref_model

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu


def _frames(seed, B=1, H=368, W=496):
    gen = torch.Generator().manual_seed(seed)
    return (torch.rand(B, 3, H, W, generator=gen) * 255).cuda(), (torch.rand(B, 3, H, W, generator=gen) * 255).cuda()


@pytest.mark.skipif(not ref_model.available(), reason="baseline/_ref not installed (tools/install_ref.py)")
@pytest.mark.parametrize("alternate", [False, True], ids=["CorrBlock", "AlternateCorrBlock"])
def test_graph_replay_equals_eager_reference_model(alternate):
    raft = ref_model.load_raft_module(dropin=True)
    torch.manual_seed(1234)
    model = raft.RAFT(ref_model.make_args(alternate_corr=alternate)).eval().cuda()
    fwd = lambda a, b: model(a, b, iters=12, test_mode=True)   # noqa: E731
    im1, im2 = _frames(1)
    step = raftcorr_b200.graphed(fwd, (im1, im2))
    for seed in (1, 2, 3):                      # the captured frames, then two new pairs
        a, b = _frames(seed)
        with torch.no_grad():
            lo_e, up_e = fwd(a, b)
        lo_g, up_g = step(a, b)
        assert torch.equal(lo_g, lo_e) and torch.equal(up_g, up_e), seed
    with pytest.raises(RuntimeError, match="captured for"):
        step(*_frames(4, H=128, W=160))


def test_graph_of_library_ops_only():
    """Without any model: build + 12 lookups + upsampling captured once, replayed for new feature maps."""
    import corr
    import upsample
    gen = torch.Generator(device="cuda").manual_seed(3)

    def fwd(f1, f2, coords, mask):
        blk = corr.CorrBlock(f1, f2, radius=4)
        acc = None
        for it in range(12):
            out = blk(coords + 0.25 * it)
            acc = out if acc is None else acc + out
        return acc, upsample.upsample_flow(coords, mask)

    mk = lambda: (torch.randn(2, 64, 24, 40, device="cuda", generator=gen), torch.randn(2, 64, 24, 40, device="cuda", generator=gen),  # noqa: E731
                  4 * torch.randn(2, 2, 24, 40, device="cuda", generator=gen) + 10, torch.randn(2, 576, 24, 40, device="cuda", generator=gen))
    step = raftcorr_b200.graphed(fwd, mk())
    for _ in range(2):
        args = mk()
        with torch.no_grad():
            want = fwd(*args)
        got = step(*args)
        assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])
