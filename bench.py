"""bench.py — RAFT frame-pairs/s at 440x1024, 12 iterations (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" = one RAFT inference (12 refinement iterations) over a batch of 8 synthetic frame
pairs per GPU: stock PyTorch/cuDNN encoders + update block (harness/raft_stock.py, random-init
weights) around THIS repo's correlation path (pytorch-raft_b200/corr.py -> libraftcorr_b200.so),
plus the library's convex-upsampling kernel (--stock-upsample switches it off).
Frame pairs shard by batch across GPUs: no collective on the data path ("scaling": "weak").

  value      whole-job frame-pairs/s with the frames already resident in HBM (CUDA events,
             barrier + synchronize on both sides, max over ranks)
  e2e        the same through the public API with HOST frames: pinned host -> device copy of
             both frames and device -> host read of the full-resolution flow inside the timed
             region, every step
  roofline   the lookup kernel (the path's dominant kernel, 12 launches per step): algorithmic
             bytes (2104 B/query on the 16-bit pyramid, 2904 B/query on the fp32
             one) / its CUDA-event duration measured live inside the timed region, against the
             measured HBM peak in MEASURED_PEAKS.json
  roofline_build   the volume build against both of its roofs (tensor pipe, HBM write)
  dropin_unchanged the UNMODIFIED reference raft.py (baseline/_ref) with this library's corr.py, and
             with the reference's own corr.py, on the same GPU and frames
  with_tc_convc1 / graph_replay_batch8
             the headline step with the lookup's consumer on the tensor cores, and
             replayed as one CUDA graph (raftcorr_b200.graphed); `value` stays the plain eager step
  cpu_baseline / --impl reference
             the reference's own CPU implementation (unmodified core/raft.py + core/corr.py from
             baseline/_ref, kind "reference"; the oracle's torch restatement in the harness model
             only when that copy is absent, kind "port") on the box's host cores, on a bounded
             sample (1 pair per step)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "pytorch-raft_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

# stdout is reserved for the one JSON line, but libraries write there too (NCCL prints its version banner with
# printf).  Keep a private duplicate of the real stdout for the result and point fd 1 at stderr for everyone else.
sys.stdout.flush()
_RESULT_FD = os.dup(1)
os.dup2(2, 1)


def emit(line: dict) -> None:
    sys.stdout.flush()
    os.write(_RESULT_FD, (json.dumps(line) + "\n").encode())


import torch  # noqa: E402

METRIC = "RAFT frame-pairs/s @440x1024 12 iters; corr lookup HBM GB/s vs peak"
UNIT = "frame-pairs/s"
IMG_H, IMG_W, ITERS, BATCH_PER_GPU = 440, 1024, 12, 8
RADIUS, LEVELS = 4, 4
# 4 levels x (2r+2)^2 window pixels read + 8 B of coordinates + 4 x (2r+1)^2 fp32 samples written:
# 2904 B per query with an fp32 pyramid, 2104 B with the 16-bit pyramid
def lookup_bytes_per_query(pixel_bytes: int) -> int:
    return LEVELS * (2 * RADIUS + 2) ** 2 * pixel_bytes + 8 + LEVELS * (2 * RADIUS + 1) ** 2 * 4


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def git_blob_sha1(path: str):
    import hashlib
    try:
        data = open(path, "rb").read()
    except OSError:
        return None
    return hashlib.sha1(b"blob %d\0" % len(data) + data).hexdigest()


def ncu_traffic_per_launch(kernel_key: str):
    """dram read+write bytes per launch of the named lookup kernel from the committed ncu --set full capture
    (profiles/lookup_traffic.json).  Every entry records the git blob hash of the kernel's source file at capture
    time; a capture of an older version of the kernel is not reported (null)."""
    path = os.path.join(ROOT, "profiles", "lookup_traffic.json")
    try:
        entry = json.load(open(path))["kernels"][kernel_key]
        if entry.get("workload") != "cfg2_b8_55x128":
            return None
        if entry.get("source_blob") != git_blob_sha1(os.path.join(PKG, "csrc", entry["source_file"])):
            return None
        return float(entry["dram_bytes_per_launch"])
    except Exception:
        return None


def measure_write_only_gbs(dev, gib: int = 2, reps: int = 6):
    """What this GPU's HBM takes when a kernel only writes (the build's situation: 1.0 GB of pyramid out, 58 MB of
    operands in): a 2 GiB fill of 4-byte elements, best of `reps` (torch's 1-byte fill kernel is itself slow, 3.9 TB/s;
    tools/probe/write_roof_probe.cu: cudaMemsetAsync 7.3 TB/s, plain 16-byte stores 6.9-7.0, TMA bulk stores 6.4)."""
    buf = torch.empty((gib << 30) // 4, dtype=torch.float32, device=dev)
    buf.zero_()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        buf.zero_()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return (gib << 30) / (best * 1e-3) / 1e9


def measure_tf32_peak(dev, seconds: float = 1.5):
    """Dense TF32 peak on this GPU, measured the way MEASURED_PEAKS.json measures bf16: torch.matmul 8192^3 with
    allow_tf32 (cuBLAS), best of 10 (burst) and back to back for `seconds` (sustained).  SURVEY.md 8(d)."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        flops = 2.0 * n ** 3
        for _ in range(3):
            a @ b
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(int(seconds * 1e3 / best), 10)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record()
        torch.cuda.synchronize()
        sustained = e0.elapsed_time(e1) / reps
        return {"tf32_tflops": flops / best / 1e9, "tf32_tflops_sustained": flops / sustained / 1e9,
                "how": f"torch.matmul fp32 8192^3 with allow_tf32: best of 10 (burst), {reps} back to back (sustained)"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None
        self.t0 = self.t1 = None   # the timed window (perf_counter); samples outside are dropped

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        for stamp, r in self.rows:
            if self.t0 is not None and not (self.t0 <= stamp <= (self.t1 or stamp)):
                continue
            try:
                sm.append(int(float(r[0])))
                mx = max(mx, int(float(r[1])))
            except (ValueError, IndexError):
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_model(device):
    from harness.raft_stock import StockRAFT
    torch.manual_seed(1234)
    return StockRAFT().eval().to(device)


def synthetic_frames(batch, seed, pin=False):
    gen = torch.Generator().manual_seed(seed)
    a = torch.rand(batch, 3, IMG_H, IMG_W, generator=gen) * 255.0
    b = torch.rand(batch, 3, IMG_H, IMG_W, generator=gen) * 255.0
    return (a.pin_memory(), b.pin_memory()) if pin else (a, b)


# ------------------------------------------------------------------------------------------------
def run_cpu_reference(steps: int, warmup: int, sample_pairs: int = 1):
    """The reference's own CPU implementation of the path on the host cores: the UNMODIFIED core/raft.py +
    core/corr.py from baseline/_ref (tools/install_ref.py; kind "reference").  Only when that copy is absent
    does the oracle's torch restatement inside the harness model stand in (kind "port")."""
    from harness import ref_model
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    im1, im2 = synthetic_frames(sample_pairs, seed=99)
    if ref_model.available():
        raft = ref_model.load_raft_module(dropin=False)
        torch.manual_seed(1234)
        ref = raft.RAFT(ref_model.make_args()).eval()
        step = lambda: ref(im1, im2, iters=ITERS, test_mode=True)  # noqa: E731
        kind, what = "reference", "the unmodified reference core/raft.py + core/corr.py (baseline/_ref) on CPU"
    else:
        from oracle import torch_restatement as tr
        model = make_model("cpu")
        factory = lambda f1, f2, r: tr.TorchCorrBlock(f1, f2, LEVELS, r)  # noqa: E731
        step = lambda: model(im1, im2, factory, iters=ITERS)  # noqa: E731
        kind, what = "port", ("oracle/torch_restatement.py (same ATen ops as core/corr.py) inside "
                              "harness/raft_stock.py; baseline/_ref not installed")
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            step()
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    total = sum(times)
    return {"value": sample_pairs * len(times) / total, "ms_per_step": 1e3 * total / len(times),
            "cores": threads, "kind": kind,
            "sample": f"{sample_pairs} frame pair(s) per step at 440x1024, 12 iterations, fp32, {len(times)} timed step(s); "
                      + what}


def measure_dropin_unchanged(dev, frames, steps: int = 5):
    """What a user of the reference gets by putting pytorch-raft_b200/ ahead of core/ on sys.path and changing nothing
    else: the UNMODIFIED reference raft.py (baseline/_ref, tools/install_ref.py) with this library's CorrBlock, same
    frames, 12 iterations, test_mode.  The same model with the reference's own corr.py is timed beside it (stock
    PyTorch on the same GPU).  Falls back to the harness model without its hooks when baseline/_ref is absent."""
    from harness import ref_model

    def timed(fn, reps, warm=2):
        for _ in range(warm):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    B = frames[0].shape[0]
    out = {}
    if ref_model.available():
        for key, dropin in (("with_this_library", True), ("reference_corr_py_on_this_gpu", False)):
            raft = ref_model.load_raft_module(dropin=dropin)
            torch.manual_seed(1234)
            net = raft.RAFT(ref_model.make_args()).eval().to(dev)
            with torch.no_grad():
                ms = timed(lambda: net(frames[0], frames[1], iters=ITERS, test_mode=True), steps if dropin else 3)
            out[key] = {"ms_per_step": ms, "frame_pairs_per_s": B / (ms * 1e-3)}
            del net
        out["model"] = "unmodified reference core/raft.py (baseline/_ref)"
    else:
        import corr as raftcorr_dropin
        model = make_model(dev)
        factory = lambda a, b, r: raftcorr_dropin.CorrBlock(a, b, num_levels=LEVELS, radius=r)  # noqa: E731
        with torch.no_grad():
            ms = timed(lambda: model(frames[0], frames[1], factory, iters=ITERS), steps)
        out["with_this_library"] = {"ms_per_step": ms, "frame_pairs_per_s": B / (ms * 1e-3)}
        out["model"] = "harness/raft_stock.py without hooks (baseline/_ref not installed)"
    out["workload"] = "configs[1]: 440x1024, 12 iterations, batch 8, drop-in only (update.py and raft.py unchanged)"
    return out


def measure_other_configs(dev, hbm_peak_gbs, model=None, frames=None):
    """Kernel-level numbers for the path's other BASELINE.json configurations (CUDA events, rank 0,
    N=1 only; a few hundred ms): configs[3] on-the-fly path at 1088x1920, configs[4] training-step
    correlation forward + backward.  Reported next to the headline, not part of it."""
    import corr as raftcorr_dropin

    def grid(B, H, W):
        ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
        return torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)

    def timed(fn, reps, warm=2):
        for _ in range(warm):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    out = {}
    gen = torch.Generator(device=dev).manual_seed(7)
    # configs[3]: AlternateCorrBlock, 136x240 fmaps, D=256, batch 4, smooth flow
    B, D, H, W = 4, 256, 136, 240
    f1 = torch.randn(B, D, H, W, device=dev, generator=gen)
    f2 = torch.randn(B, D, H, W, device=dev, generator=gen)
    low = 8 * torch.randn(B, 2, H // 8, W // 8, device=dev, generator=gen)
    c = grid(B, H, W) + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
    with torch.no_grad():
        blk = raftcorr_dropin.AlternateCorrBlock(f1, f2)
        ms = timed(lambda: blk(c), reps=10)
    N = H * W
    T = sum((H >> l) * (W >> l) for l in range(4))
    compulsory = B * (N * D * 4 + T * D * 4 + 8 * N + 1296 * N)          # 120.4 MB per pair
    useful_flops = B * N * 4 * 100 * 2 * D
    physical = B * (N * D * 2 + T * D * 2 + 8 * N + 1296 * N)            # the kernel reads bf16 copies of the fmaps
    out["alt_cfg4"] = {"workload": "AlternateCorrBlock 1088x1920 (136x240 fmaps), batch 4, D=256, smooth flow",
                       "ms_per_iter": ms, "compulsory_gb_s": compulsory / ms / 1e6,
                       "frac_hbm_at_compulsory_bytes": compulsory / ms / 1e6 / hbm_peak_gbs,
                       "frac_hbm_at_physical_bf16_bytes": physical / ms / 1e6 / hbm_peak_gbs,
                       "useful_tflops": useful_flops / ms / 1e9, "binding": "sampler issue slots (tensor pipe 27 % active)"}
    del f1, f2, blk
    # configs[4]: CorrBlock forward + backward of 12 lookups at 368x496, batch 12
    B, D, H, W = 12, 256, 46, 62
    f1 = torch.randn(B, D, H, W, device=dev, generator=gen, requires_grad=True)
    f2 = torch.randn(B, D, H, W, device=dev, generator=gen, requires_grad=True)
    c = grid(B, H, W) + 4 * torch.randn(B, 2, H, W, device=dev, generator=gen)

    def train_step():
        blk = raftcorr_dropin.CorrBlock(f1, f2)
        loss = 0
        for _ in range(ITERS):
            loss = loss + blk(c).sum()
        loss.backward()
        f1.grad = f2.grad = None

    out["train_cfg5"] = {"workload": "CorrBlock build + 12 lookups, forward and backward, 368x496 (46x62 fmaps), batch 12",
                         "ms_per_step": timed(train_step, reps=3, warm=1)}
    del f1, f2
    # configs[4] as written, on this one GPU: the whole training step of the unmodified reference model (forward, loss,
    # backward, clip, AdamW step; harness/train_step.py) with this library's corr.py and with the reference's own
    from harness import ref_model as _rm, train_step as _ts
    if _rm.available():
        def whole_step(dropin):
            tr = _ts.Trainer(dev, dropin=dropin)
            batch = _ts.synthetic_batch(_ts.BATCH_PER_GPU, dev, seed=500)
            ms = timed(lambda: tr.step(*batch), reps=3, warm=2)
            del tr
            torch.cuda.empty_cache()
            return ms
        ours_ms, theirs_ms = whole_step(True), whole_step(False)
        out["train_step_cfg5"] = {
            "workload": "unmodified reference RAFT, whole training step, 368x496, batch 12, 12 iterations, 1 GPU "
                        "(python bench.py --workload train under torchrun gives the N-GPU DDP line)",
            "ms_per_step": ours_ms, "frame_pairs_per_s": _ts.BATCH_PER_GPU / (ours_ms * 1e-3),
            "corr_fwd_bwd_share_of_step": out["train_cfg5"]["ms_per_step"] / ours_ms,
            "same_step_with_reference_corr_py_ms": theirs_ms}
    # configs[2]: KITTI shape, 47x156 fmaps, batch 16, 24 iterations: the correlation path alone
    B, D, H, W = 16, 256, 47, 156
    f1 = torch.randn(B, D, H, W, device=dev, generator=gen)
    f2 = torch.randn(B, D, H, W, device=dev, generator=gen)
    c = grid(B, H, W) + 4 * torch.randn(B, 2, H, W, device=dev, generator=gen)

    def kitti_path():
        with torch.no_grad():
            blk = raftcorr_dropin.CorrBlock(f1, f2)
            for _ in range(24):
                blk(c)

    out["corr_path_cfg3"] = {"workload": "CorrBlock build + 24 lookups, KITTI 376x1248 (47x156 fmaps), batch 16",
                             "ms_per_step": timed(kitti_path, reps=3, warm=1)}
    del f1, f2
    # the step after the path: convex upsampling at configs[1] size, this library against the reference's ops
    import upsample as raftcorr_upsample
    from oracle import torch_restatement as tr
    B, H, W = BATCH_PER_GPU, IMG_H // 8, IMG_W // 8
    masks = [torch.randn(B, 576, H, W, device=dev, generator=gen) for _ in range(4)]   # 4 x 130 MB > L2
    flow = torch.randn(B, 2, H, W, device=dev, generator=gen)
    turn = {"i": 0}

    def next_mask():
        turn["i"] += 1
        return masks[turn["i"] & 3]

    with torch.no_grad():
        ours = timed(lambda: raftcorr_upsample.upsample_flow(flow, next_mask()), reps=20)
        stock = timed(lambda: tr.upsample_flow(flow, next_mask()), reps=5)
    up_bytes = B * H * W * (2304 + 8 + 512)
    out["upsample_flow"] = {"workload": "RAFT.upsample_flow at 55x128, batch 8 (SURVEY.md 8f-4)", "us": ours * 1e3,
                            "algorithmic_gb_s": up_bytes / ours / 1e6, "frac_hbm": up_bytes / ours / 1e6 / hbm_peak_gbs,
                            "same_ops_as_reference_on_this_gpu_us": stock * 1e3}
    del masks
    # the reference's own op sequence on this GPU through stock PyTorch (ATen + cuBLAS; SURVEY.md 8d "additional
    # honest baseline"): same model, same frames, only the correlation block and the upsampling differ
    if model is not None and frames is not None:
        def stock_step():
            with torch.no_grad():
                model(frames[0], frames[1], lambda a, b, r: tr.TorchCorrBlock(a, b, LEVELS, r), iters=ITERS)

        ms = timed(stock_step, reps=3, warm=1)
        # batch 1 (what evaluate.py / demo.py run): the loop is launch-bound, so the same forward is also timed as
        # one CUDA graph replay through the product helper raftcorr_b200.graphed — with the
        # unmodified reference model when baseline/_ref is installed, else with the harness model
        import raftcorr_b200
        from harness import ref_model as _rm2
        one = (frames[0][:1].contiguous(), frames[1][:1].contiguous())
        if _rm2.available():
            torch.manual_seed(1234)
            net = _rm2.load_raft_module(dropin=True).RAFT(_rm2.make_args()).eval().to(dev)
            b1_fwd = lambda a, b: net(a, b, iters=ITERS, test_mode=True)[1]  # noqa: E731
            which = "unmodified reference raft.py + this library's corr.py"
        else:
            factory = lambda a, b, r: raftcorr_dropin.CorrBlock(a, b, num_levels=LEVELS, radius=r)  # noqa: E731
            b1_fwd = lambda a, b: model(a, b, factory, iters=ITERS)[1]  # noqa: E731
            which = "harness model + this library's corr.py"
        with torch.no_grad():
            eager_ms = timed(lambda: b1_fwd(*one), reps=5, warm=2)
        step = raftcorr_b200.graphed(b1_fwd, one)
        graph_ms = timed(lambda: step(*one), reps=10, warm=2)
        out["latency_batch1"] = {"workload": "configs[1] shape at batch 1, 12 iterations, " + which,
                                 "eager_ms": eager_ms, "cuda_graph_replay_ms": graph_ms,
                                 "how": "raftcorr_b200.graphed(forward, (image1, image2))"}
        del step
        # the headline step (batch 8, harness model, library CorrBlock + upsampler) replayed as one CUDA graph: what the
        # launch gaps of ~1300 eager launches per step cost (`value` stays the eager figure)
        import upsample as _up
        fac8 = lambda a, b, r: raftcorr_dropin.CorrBlock(a, b, num_levels=LEVELS, radius=r)  # noqa: E731

        def b8_fwd(a, b):
            with torch.no_grad():
                return model(a, b, fac8, iters=ITERS, upsampler=_up.upsample_flow)[1]
        eager8 = b8_fwd(*frames).clone()
        step8 = raftcorr_b200.graphed(b8_fwd, frames)
        same = bool(torch.equal(step8(*frames), eager8))
        g8 = timed(lambda: step8(*frames), reps=10, warm=2)
        out["graph_replay_batch8"] = {"workload": "the headline step (batch 8) as one CUDA graph replay, raftcorr_b200.graphed",
                                      "ms_per_step": g8, "frame_pairs_per_s": BATCH_PER_GPU / (g8 * 1e-3),
                                      "bit_identical_to_eager": same}
        del step8, eager8
        out["reference_ops_on_this_gpu"] = {
            "workload": "configs[1] with the reference's CorrBlock / upsample_flow op sequence (stock PyTorch) on the B200",
            "ms_per_step": ms, "frame_pairs_per_s": BATCH_PER_GPU / (ms * 1e-3)}
    return out


def run_train_workload(args, rank: int, world: int, local_rank: int):
    """BASELINE.json configs[4]: whole training step (harness/train_step.py), one process per GPU, NCCL all-reduce
    through DistributedDataParallel.  Not the headline metric: printed as its own JSON line."""
    from harness import ref_model, train_step
    if not ref_model.available():
        if rank == 0:
            emit({"workload": "train_cfg5", "unavailable": "baseline/_ref not installed (tools/install_ref.py)"})
        return
    distributed = world > 1
    if distributed:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank if distributed else 0)
    torch.cuda.set_device(dev)
    steps, warmup = args.steps, max(args.warmup, 3)

    def measure(dropin: bool):
        tr = train_step.Trainer(dev, dropin=dropin, distributed=distributed, local_rank=local_rank,
                                precision=args.precision)
        batch = train_step.synthetic_batch(train_step.BATCH_PER_GPU, dev, seed=500 + rank)
        for _ in range(warmup):
            tr.step(*batch)
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(steps):
            loss, _ = tr.step(*batch)
        t1.record()
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        if distributed:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        peak_gb = torch.cuda.max_memory_allocated(dev) / 2 ** 30
        del tr
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats(dev)
        return float(ms.item()) / steps, float(loss.item()), peak_gb

    ours_ms, ours_loss, ours_gb = measure(True)
    ref_ms = ref_loss = ref_gb = None
    if not args.skip_cpu_baseline:       # the same step with the reference's own corr.py (ATen ops) on the same GPUs
        ref_ms, ref_loss, ref_gb = measure(False)
    if rank == 0:
        pairs = train_step.BATCH_PER_GPU * world
        emit({"workload": "train_cfg5", "metric": "RAFT training step frame-pairs/s @368x496 12 iters (BASELINE.json configs[4])",
              "value": pairs / (ours_ms * 1e-3), "unit": "frame-pairs/s", "n_gpus": world, "steps": steps, "warmup": warmup,
              "ms_per_step": ours_ms, "higher_is_better": True, "scaling": "weak", "dtype": args.precision or "bf16",
              "data": "synthetic",
              "config": {"workload": "unmodified reference RAFT (baseline/_ref) forward + sequence_loss + backward + clip + "
                                     "AdamW/OneCycle step, 368x496, batch 12 per GPU, 12 iterations, fp32 networks",
                         "corr": "this library's CorrBlock (fp32 pyramid for gradients) and autograd kernels",
                         "gradient_allreduce": "DistributedDataParallel over NCCL" if distributed else "single GPU, none"},
              "loss_after_timed_steps": ours_loss, "peak_memory_gib": ours_gb,
              "same_step_with_reference_corr_py": None if ref_ms is None else
              {"ms_per_step": ref_ms, "frame_pairs_per_s": pairs / (ref_ms * 1e-3), "loss_after_timed_steps": ref_loss,
               "peak_memory_gib": ref_gb}})
    if distributed:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="infer", choices=["infer", "train"],
                    help="infer: the headline metric (configs[1]); train: configs[4], whole training step under DDP")
    ap.add_argument("--precision", default=None, help="bf16 (default) | tf32 | fp32 for the volume build")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--tc-convc1", action="store_true",
                    help="run convc1 + bias + ReLU (update.py:83,90) as the library's tensor-core 1x1 convolution on a "
                         "channels-last lookup (needs the fp32 pyramid and an edited update.py) instead of stock torch ops")
    ap.add_argument("--pyramid", default="auto", choices=["auto", "fp32", "bf16"],
                    help="pyramid storage: auto = blocked bf16 whenever the build arithmetic is bf16")
    ap.add_argument("--stock-upsample", action="store_true",
                    help="run RAFT.upsample_flow as the reference's torch op sequence instead of this library's kernel")
    ap.add_argument("--cudnn-benchmark", action="store_true", help="let cuDNN autotune the stock convolutions")
    ap.add_argument("--channels-last", action="store_true",
                    help="run the stock model in torch.channels_last and take the lookup in the same format")
    args = ap.parse_args()
    warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    base_config = {"workload": "RAFT inference, Sintel shape 440x1024 (55x128 fmaps), 12 iters, "
                               f"batch {BATCH_PER_GPU} per GPU, random-init weights",
                   "image": [IMG_H, IMG_W], "iters": ITERS, "batch_per_gpu": BATCH_PER_GPU,
                   "corr": "CorrBlock 4 levels r=4", "sharding": f"batch x{max(world, args.gpus)}, no collective"}

    if args.impl == "reference":
        if rank != 0:
            return
        ref = run_cpu_reference(args.steps, warmup)
        line = {"impl": "reference", "metric": METRIC, "value": ref["value"], "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": warmup, "ms_per_step": ref["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict(base_config, batch_per_step=1, where="host CPU"),
                "cpu_baseline": {"value": ref["value"], "unit": UNIT, "cores": ref["cores"], "kind": ref["kind"],
                                 "sample": ref["sample"]},
                "e2e": {"value": ref["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return

    assert torch.cuda.is_available(), "bench.py (impl ours) needs CUDA devices; there is no CPU fallback"
    if args.workload == "train":
        return run_train_workload(args, rank, world, local_rank)
    import corr as raftcorr_dropin  # the drop-in module; raises if libraftcorr_b200.so is missing

    distributed = world > 1
    if distributed:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank if distributed else 0)
    torch.cuda.set_device(dev)

    torch.backends.cudnn.benchmark = bool(args.cudnn_benchmark)
    model = make_model(dev)
    host1, host2 = synthetic_frames(BATCH_PER_GPU, seed=100 + rank, pin=True)
    dev1, dev2 = host1.to(dev), host2.to(dev)
    if args.channels_last:
        model = model.to(memory_format=torch.channels_last)
        dev1, dev2 = dev1.contiguous(memory_format=torch.channels_last), dev2.contiguous(memory_format=torch.channels_last)
    host_flow = torch.empty(BATCH_PER_GPU, 2, IMG_H, IMG_W).pin_memory()
    args.stock_convc1 = not args.tc_convc1
    nhwc_lookup = args.channels_last    # (--tc-convc1 reads the plain NCHW lookup: raftcorr_conv1x1_nchw_fwd)
    storage = "fp32" if nhwc_lookup else args.pyramid          # the channels-last lookup reads the fp32 pyramid
    from raftcorr_b200 import ops
    storage16 = ops.resolve_storage(storage, ops.resolve_precision(args.precision), 256, LEVELS, RADIUS) == "bf16"
    factory = lambda f1, f2, r: raftcorr_dropin.CorrBlock(f1, f2, num_levels=LEVELS, radius=r,  # noqa: E731
                                                          precision=args.precision, channels_last=nhwc_lookup,
                                                          storage=storage)
    convc1_fn = None if args.stock_convc1 else raftcorr_dropin.conv1x1_relu

    lookup_events = []

    def timed_lookup(corr_fn, coords):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = corr_fn(coords)
        b.record()
        lookup_events.append((a, b))
        return out

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    upsampler = None
    if not args.stock_upsample:
        import upsample as raftcorr_upsample
        upsampler = raftcorr_upsample.upsample_flow

    def step_resident(hook=None):
        with torch.no_grad():
            return model(dev1, dev2, factory, iters=ITERS, lookup_hook=hook, upsampler=upsampler, convc1_fn=convc1_fn)[1]

    # End to end: every step copies its frames from pinned host memory and returns its flow to pinned host
    # memory.  The copies run on a second stream so that step i+1's frames travel while step i computes and
    # step i's flow travels while step i+1 computes (two device frame buffers, two host result buffers); the
    # host waits for step i-1's result before it leaves step i, and timed()'s final synchronize drains the rest.
    copy_stream = torch.cuda.Stream(dev)
    frame_slots = [(torch.empty_like(dev1), torch.empty_like(dev2)) for _ in range(2)]
    host_flows = [host_flow, torch.empty_like(host_flow).pin_memory()]
    ev_ready = [torch.cuda.Event() for _ in range(2)]    # frames of this slot are on the device
    ev_free = [torch.cuda.Event() for _ in range(2)]     # the model has finished reading this slot
    ev_out = [torch.cuda.Event() for _ in range(2)]      # this slot's flow is in host memory
    e2e_state = {"i": 0, "primed": False}

    def prefetch_frames(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_free[slot])
            frame_slots[slot][0].copy_(host1, non_blocking=True)
            frame_slots[slot][1].copy_(host2, non_blocking=True)
            ev_ready[slot].record(copy_stream)

    def step_e2e():
        i = e2e_state["i"]
        slot = i & 1
        main = torch.cuda.current_stream()
        if not e2e_state["primed"]:
            prefetch_frames(slot)
            e2e_state["primed"] = True
        prefetch_frames(slot ^ 1)
        with torch.no_grad():
            main.wait_event(ev_ready[slot])
            flow = model(frame_slots[slot][0], frame_slots[slot][1], factory, iters=ITERS, upsampler=upsampler,
                         convc1_fn=convc1_fn)[1]
            ev_free[slot].record(main)
            flow.record_stream(copy_stream)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ev_free[slot])
                host_flows[slot].copy_(flow, non_blocking=True)
                ev_out[slot].record(copy_stream)
        if i > 0:
            ev_out[slot ^ 1].synchronize()
        e2e_state["i"] = i + 1
        return host_flows[slot]

    for _ in range(warmup):
        step_resident()
    step_e2e()

    def timed(fn, steps, drain=None):
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(steps):
            fn()
        if drain is not None:
            drain()              # work queued on other streams completes inside the timed region
        t1.record()
        barrier()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        if distributed:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(dev.index or 0)
    if rank == 0:
        sampler.start()          # nvidia-smi needs a moment to start: launch it before warm-up ends
        for _ in range(2):
            step_resident()
        sampler.t0 = time.perf_counter()
    profile_range = bool(os.environ.get("BENCH_CUDA_PROFILER_RANGE"))  # ncu --profile-from-start off
    if profile_range:
        torch.cuda.profiler.start()
    ms_total = timed(lambda: step_resident(timed_lookup), args.steps)
    if profile_range:
        torch.cuda.profiler.stop()
    if rank == 0:
        sampler.t1 = time.perf_counter()
    clocks = sampler.stop() if rank == 0 else None
    # the last step's device -> host copy runs on copy_stream: make the timing stream wait for it before t1
    ms_e2e = timed(step_e2e, args.steps, drain=lambda: torch.cuda.current_stream().wait_stream(copy_stream))

    # the same step with the lookup's consumer (convc1 + bias + ReLU) on the tensor cores, reading the NCHW lookup as an
    # MN-major operand (needs a two-line edit of update.py, hence not the headline)
    ms_tcconv = None
    if args.stock_convc1 and not nhwc_lookup:
        def step_tcconv():
            with torch.no_grad():
                return model(dev1, dev2, factory, iters=ITERS, upsampler=upsampler, convc1_fn=raftcorr_dropin.conv1x1_relu)[1]
        for _ in range(2):
            step_tcconv()
        ms_tcconv = timed(step_tcconv, args.steps) / args.steps

    # kernel shares: the build (pack + tcgen05 GEMM with the pooled epilogue) once per step, timed on its own with the
    # operands rotated so that no launch finds its inputs or outputs in L2
    fms = [(torch.randn(BATCH_PER_GPU, 256, IMG_H // 8, IMG_W // 8, device=dev),
            torch.randn(BATCH_PER_GPU, 256, IMG_H // 8, IMG_W // 8, device=dev)) for _ in range(2)]
    for i in range(2):
        ops.build_pyramid(*fms[i], LEVELS, args.precision, storage=storage, radius=RADIUS)
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b0.record()
    for i in range(6):
        ops.build_pyramid(*fms[i & 1], LEVELS, args.precision, storage=storage, radius=RADIUS)
    b1.record()
    torch.cuda.synchronize()
    build_ms = b0.elapsed_time(b1) / 6
    del fms
    write_only_gbs = measure_write_only_gbs(dev)

    if rank != 0:
        if distributed:
            dist.destroy_process_group()
        return

    n_gpus = world if distributed else 1
    pairs = BATCH_PER_GPU * n_gpus * args.steps
    lookup_ms = [a.elapsed_time(b) for a, b in lookup_events]
    lookup_avg_ms = sum(lookup_ms) / max(len(lookup_ms), 1)
    queries = BATCH_PER_GPU * (IMG_H // 8) * (IMG_W // 8)
    peak, peak_src = measured_peaks()
    bytes_per_query = lookup_bytes_per_query(2 if storage16 else 4)
    achieved = queries * bytes_per_query / (lookup_avg_ms * 1e-3) / 1e9
    if storage16:
        lookup_kernel, traffic_key = "lookup16_kernel (bf16 pyramid in 64-byte blocks, NCHW output)", "lookup16_kernel"
    elif nhwc_lookup:
        lookup_kernel, traffic_key = "lookup_fwd_tma_kernel (fp32 pyramid, channels-last output)", "lookup_fwd_tma_kernel_nhwc"
    else:
        lookup_kernel, traffic_key = "lookup_fwd_tma_kernel (fp32 pyramid, NCHW output)", "lookup_fwd_tma_kernel"
    # the build against both of its roofs: 2 N^2 D flops per pair on the tensor pipe, the pyramid
    # written once to HBM; whichever is closer to its peak binds
    peaks_all = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    Nq = (IMG_H // 8) * (IMG_W // 8)
    T = sum(((IMG_H // 8) >> l) * ((IMG_W // 8) >> l) for l in range(LEVELS))
    build_flops = 2.0 * BATCH_PER_GPU * Nq * Nq * 256
    build_bytes = BATCH_PER_GPU * Nq * (T * (2 if storage16 else 4) + 2 * 256 * 4)
    tf_sust = float(peaks_all.get("bf16_tflops_sustained", 1400.0))
    tf_burst = float(peaks_all.get("bf16_tflops", 1590.0))
    build_tflops = build_flops / (build_ms * 1e-3) / 1e12
    build_gbs = build_bytes / (build_ms * 1e-3) / 1e9
    roofline_build = {
        "kernel": "pack2_bf16_kernel + build_tc_kernel (tcgen05 M128 N256, " + ("bf16 blocked" if storage16 else "fp32") + " pyramid epilogue)",
        "ms": build_ms, "flops": build_flops, "algorithmic_bytes": build_bytes,
        "tensor": {"achieved_tflops": build_tflops, "peak_sustained": tf_sust, "frac_sustained": build_tflops / tf_sust,
                   "peak_burst": tf_burst, "frac_burst": build_tflops / tf_burst,
                   "which": "bf16 dense, cuBLAS 8192^3 (MEASURED_PEAKS.json)" if peaks_all else "fallback"},
        "hbm": {"achieved_gbs": build_gbs, "peak": peak, "frac": build_gbs / peak},
        "hbm_write_only": {"peak": write_only_gbs, "frac": build_gbs / write_only_gbs,
                           "how": "2 GiB fill of 4-byte elements on this GPU, best of 6: what a kernel that only writes can reach"},
        "bound": "neither roof: epilogue-limited (DESIGN.md 3.1)" if max(build_tflops / tf_sust, build_gbs / write_only_gbs) < 0.7
                 else ("tensor" if build_tflops / tf_sust >= build_gbs / write_only_gbs else "hbm-write"),
    }
    h2d = 2 * BATCH_PER_GPU * 3 * IMG_H * IMG_W * 4
    d2h = BATCH_PER_GPU * 2 * IMG_H * IMG_W * 4

    line = {
        "metric": METRIC, "value": pairs / (ms_total * 1e-3), "unit": UNIT, "n_gpus": n_gpus,
        "steps": args.steps, "warmup": warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if args.precision in (None, "bf16") else args.precision,
        "data": "synthetic",
        "config": dict(base_config, l2="inputs exceed L2 (2.1 GB pyramid + 115 MB frames per step)",
                       build_precision=args.precision or "bf16", encoders="stock PyTorch/cuDNN fp32",
                       pyramid="bf16 in 4x8-pixel 64-byte blocks (raftcorr_level16)" if storage16 else "fp32 rows (raftcorr_level)",
                       convc1="stock torch conv2d + relu on an NCHW lookup (update.py unchanged)" if args.stock_convc1 else
                       "raftcorr_conv1x1_nchw_fwd (tf32 tcgen05, MN-major activations, bias + ReLU epilogue) on the NCHW lookup",
                       upsample="stock torch ops (raft.py:72-83)" if args.stock_upsample else
                       "raftcorr_upsample_flow_fwd, every iteration like the reference",
                       memory_format="channels_last" if args.channels_last else "contiguous (NCHW, as the reference)"),
        "e2e": {"value": pairs / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
                "copies": "pinned host <-> device on a second stream, double-buffered, overlapping the neighbouring steps"},
        # pack + tcgen05 build + 12 lookups (+ 12 tensor-core 1x1 convolutions, + 12 convex upsamplings) per step
        "gpu_launches": (2 + ITERS * (1 + (not args.stock_upsample) + (not args.stock_convc1))) * args.steps * n_gpus,
        "roofline": {"kernel": lookup_kernel,
                     "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic_per_launch(traffic_key),
                     "peak_source": peak_src, "algorithmic_bytes_per_query": bytes_per_query,
                     "algorithmic_bytes_per_launch": queries * bytes_per_query,
                     "launch_us": lookup_avg_ms * 1e3, "launches_timed": len(lookup_ms),
                     "launch_us_by_iteration": [round(1e3 * sum(lookup_ms[i::ITERS]) / max(len(lookup_ms[i::ITERS]), 1), 2)
                                                for i in range(ITERS)],
                     "note": "launch_us averages all 12 launches of a step, including iteration 0 whose integer "
                             "coordinates take the kernel's slower floor-flip path"},
        "roofline_build": roofline_build,
        "kernels": {"lookup_us": lookup_avg_ms * 1e3, "build_ms": build_ms,
                    "corr_path_ms_per_step": build_ms + ITERS * lookup_avg_ms,
                    "corr_path_share_of_step": (build_ms + ITERS * lookup_avg_ms) / (ms_total / args.steps)},
        "clocks": clocks,
    }
    if ms_tcconv is not None:
        line["with_tc_convc1"] = {"ms_per_step": ms_tcconv, "frame_pairs_per_s": BATCH_PER_GPU * n_gpus / (ms_tcconv * 1e-3),
                                  "convc1": "raftcorr_conv1x1_nchw_fwd on the NCHW lookup (update.py edited as in INTEGRATION.md 2c)"}
    if n_gpus == 1:
        line["other_configs"] = measure_other_configs(dev, peak, model, (dev1, dev2))
        if "latency_batch1" in line["other_configs"]:
            line["latency_batch1"] = line["other_configs"]["latency_batch1"]
        if "graph_replay_batch8" in line["other_configs"]:
            line["graph_replay_batch8"] = line["other_configs"].pop("graph_replay_batch8")
        line["dropin_unchanged"] = measure_dropin_unchanged(dev, (dev1, dev2))
        line["measured_here"] = measure_tf32_peak(dev)      # last: a seconds-long GEMM burn pulls the clocks down
    if not args.skip_cpu_baseline and n_gpus == 1:
        ref = run_cpu_reference(steps=1, warmup=1)
        line["cpu_baseline"] = {"value": ref["value"], "unit": UNIT, "cores": ref["cores"], "kind": ref["kind"],
                                "sample": ref["sample"]}
    emit(line)
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
