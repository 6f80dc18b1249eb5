"""Drop-in hygiene: the library is called from one Python thread per GPU under
nn.DataParallel and must honour the caller's stream; it must also be capturable in a CUDA graph
by a caller who wants to remove launch overhead from RAFT's refinement loop."""
import threading

import numpy as np
import pytest
import torch

import oracle
from raftcorr_b200 import ops
from helpers import coords_for, pyramid_from_numpy

pytestmark = pytest.mark.gpu


def test_concurrent_threads_and_streams_same_device():
    gen = torch.Generator().manual_seed(13)
    cases = []
    for k, (B, H, W) in enumerate([(1, 24, 40), (2, 16, 36), (1, 31, 33), (1, 24, 40)]):
        levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen).numpy() for l in range(4)]
        c = coords_for("sigma4", B, H, W, gen)
        cases.append((levels, c, oracle.lookup(levels, c.numpy(), 4), (B, H, W)))
    results, errors = [None] * len(cases), []

    def worker(idx):
        try:
            levels, c, _, (B, H, W) = cases[idx]
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                pyr = pyramid_from_numpy(levels, B, H, W)
                cd = c.cuda()
                for _ in range(25):          # many calls: exercises the tensor-map cache under contention
                    out = ops.lookup(pyr, cd, 4, "aten_cpu")
                stream.synchronize()
                results[idx] = out.cpu().numpy()
        except Exception as exc:   # pragma: no cover
            errors.append(exc)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(len(cases))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for got, (_, _, want, _) in zip(results, cases):
        assert np.array_equal(got, want)


def test_build_and_lookup_are_cuda_graph_capturable():
    gen = torch.Generator(device="cuda").manual_seed(3)
    f1 = torch.randn(2, 64, 24, 40, device="cuda", generator=gen)
    f2 = torch.randn(2, 64, 24, 40, device="cuda", generator=gen)
    c = coords_for("sigma4", 2, 24, 40, torch.Generator().manual_seed(1)).cuda()
    eager = ops.lookup(ops.build_pyramid(f1, f2, 4, "bf16"), c, 4)   # also warms one-time set-up
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            pyr = ops.build_pyramid(f1, f2, 4, "bf16")
            out = ops.lookup(pyr, c, 4)
    torch.cuda.current_stream().wait_stream(side)
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)
    f1.mul_(2.0)               # the graph re-reads its inputs: outputs follow them
    g.replay()
    torch.cuda.synchronize()
    assert torch.allclose(out, 2.0 * eager, rtol=2e-2, atol=1e-3)


def test_whole_refinement_loop_replays_as_one_cuda_graph():
    """a full 12-iteration RAFT forward (stock encoders + this library's build, lookups and
    upsampling) captured once and replayed gives the eager result; the on-the-fly block and upflow8 capture too."""
    import corr
    import upsample
    from harness.raft_stock import StockRAFT
    torch.manual_seed(11)
    model = StockRAFT().eval().cuda()
    gen = torch.Generator().manual_seed(2)
    im1 = (torch.rand(1, 3, 192, 256, generator=gen) * 255).cuda()
    im2 = (torch.rand(1, 3, 192, 256, generator=gen) * 255).cuda()

    def run(factory):
        with torch.no_grad():
            lo, up = model(im1, im2, factory, iters=12, upsampler=upsample.upsample_flow)
            return lo, up, upsample.upflow8(lo)

    for factory in (lambda a, b, r: corr.CorrBlock(a, b, radius=r),
                    lambda a, b, r: corr.AlternateCorrBlock(a, b, radius=r)):
        eager = [t.clone() for t in run(factory)]       # also warms cuDNN and the one-time kernel set-up
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            run(factory)
            with torch.cuda.graph(g, stream=side):
                outs = run(factory)
        torch.cuda.current_stream().wait_stream(side)
        for t in outs:
            t.zero_()
        g.replay()
        torch.cuda.synchronize()
        for got, want in zip(outs, eager):
            assert torch.allclose(got, want, rtol=1e-4, atol=1e-4), (got - want).abs().max().item()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_data_parallel_style_threads_on_two_devices():
    """nn.DataParallel (train.py:138) calls the block from one thread per GPU: every entry point must work on
    the tensors' device whatever the thread's current device is, with identical results on both devices."""
    import corr
    import upsample
    gen = torch.Generator().manual_seed(21)
    B, D, H, W = 2, 64, 24, 40
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    c = coords_for("sigma4", B, H, W, gen)
    flow = torch.randn(B, 2, H, W, generator=gen)
    mask = torch.randn(B, 576, H, W, generator=gen)
    gout = torch.randn(B, 324, H, W, generator=gen)
    results, errors = {}, []

    def worker(dev, current):
        try:
            torch.cuda.set_device(current)          # deliberately not always the tensors' device
            d = torch.device("cuda", dev)
            a, b = f1.to(d).requires_grad_(True), f2.to(d).requires_grad_(True)
            out = corr.CorrBlock(a, b)(c.to(d))
            out.backward(gout.to(d))
            with torch.no_grad():
                alt = corr.AlternateCorrBlock(f1.to(d), f2.to(d))(c.to(d))
                up = upsample.upsample_flow(flow.to(d), mask.to(d))
            torch.cuda.synchronize(d)
            results[dev] = [t.detach().cpu() for t in (out, a.grad, b.grad, alt, up)]
        except Exception as exc:   # pragma: no cover
            errors.append((dev, exc))

    threads = [threading.Thread(target=worker, args=(0, 1)), threading.Thread(target=worker, args=(1, 0))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for x, y in zip(results[0], results[1]):
        assert torch.allclose(x, y, rtol=1e-3, atol=1e-4 * y.abs().max().item())
    assert torch.equal(results[0][0], results[1][0]) and torch.equal(results[0][4], results[1][4])
