"""Host logic of the multi-GPU training step (harness/train_step.py, bench.py --workload train) on CPU: two gloo
ranks wrap the unmodified reference model in DistributedDataParallel, take one step on different batches and must
end with identical parameters (the gradient all-reduce is the only exchange; BASELINE.json configs[4]).  The
correlation block on this CPU path is the reference's own corr.py — this library has no CPU fallback."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from harness import train_step
    tr = train_step.Trainer("cpu", dropin=False, distributed=True, iters=2)
    batch = train_step.synthetic_batch(1, "cpu", seed=10 + rank, size=(128, 160))   # 16x20 fmaps: every pyramid level at least 2x2 (a 1-pixel level divides by zero in the reference)
    loss, metrics = tr.step(*batch)
    flat = torch.cat([p.detach().reshape(-1) for p in tr.net.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    if rank == 0:
        out.put((float(loss), bool(torch.isfinite(flat).all()), bool(torch.equal(gathered[0], gathered[1])), metrics["epe"]))
    dist.destroy_process_group()


def test_two_rank_training_step_keeps_replicas_identical():
    from harness import ref_model
    if not ref_model.available():
        pytest.skip("reference tree not present")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    loss, finite, same, epe = out.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert finite and loss > 0 and epe > 0
    assert same, "replicas diverged after one DistributedDataParallel step"
