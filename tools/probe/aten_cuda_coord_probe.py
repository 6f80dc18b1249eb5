"""Which fp32 op sequence does ATen-CUDA apply to 2*x/(W-1)-1 ?  (GPU box only)"""
import numpy as np, torch
f = np.float32
ps = np.concatenate([np.arange(-6, 140, dtype=np.float32), np.arange(-6, 140, dtype=np.float32) + f(0.5),
                     (np.random.RandomState(0).randn(4000) * 40).astype(np.float32)])
for s in (64, 128, 27, 55, 16, 6, 32, 13, 156, 47):
    sm1 = f(s - 1)
    x = torch.from_numpy(ps).cuda()
    got = (2 * x / (s - 1) - 1).cpu().numpy()
    inv = f(1) / sm1
    cands = {
        "div": (f(2) * ps) / sm1 - f(1),
        "mul_inv": (f(2) * ps) * inv - f(1),
        "x*(2/sm1)": ps * (f(2) / sm1) - f(1),
        "fma(2x,inv,-1)": (np.float64(f(2) * ps) * np.float64(inv) - 1.0).astype(np.float32),
    }
    print(s, {k: int((v != got).sum()) for k, v in cands.items()})
    # separate steps
    t1 = (2 * x); t2 = (t1 / (s - 1)).cpu().numpy()
    print("    div step: vs true div", int((t2 != (f(2) * ps) / sm1).sum()), "vs mul inv", int((t2 != (f(2) * ps) * inv).sum()))
# grid_sample un-normalise probe: ramp image, sample at normalised coords, compare floor/zero pattern
for s in (64, 27):
    img = torch.zeros(1, 1, 1, s, device="cuda"); img[0, 0, 0, 0] = 1.0   # one-hot at x=0
    xn = torch.tensor([-1.0317460298538208, -1.0317461490631104, -1.031746, -1.0], device="cuda")
    grid = torch.stack([xn, torch.full_like(xn, -1.0)], -1).view(1, 1, -1, 2)
    out = torch.nn.functional.grid_sample(img.expand(1, 1, 2, s).contiguous(), grid, align_corners=True)
    print("grid_sample one-hot@0 size", s, out.flatten().tolist())
