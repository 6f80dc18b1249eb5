"""Stock-PyTorch caller of the correlation path: RAFT's "basic" model (reference core/raft.py,
core/extractor.py:118-192, core/update.py:79-136) restated with torch.nn only.

Everything here is library code (cuDNN convolutions, ATen elementwise ops); none of it is the
optimised path.  It exists so that (a) bench.py can time whole-model frame-pairs/s on the GPU
box, with either this repo's CorrBlock or the oracle's CPU restatement plugged in, and (b) the
12-iteration end-point-error test has a caller.  tests/test_harness.py loads one state_dict into
this model and into the unmodified reference model (when /root/reference is present) and checks
that the flows agree, which pins the harness.

The correlation block is injected: ``corr_factory(fmap1, fmap2, radius) -> fn(coords)`` —
exactly how core/raft.py:104-107,124 uses CorrBlock.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _norm(kind: str, ch: int) -> nn.Module:
    if kind == "instance":
        return nn.InstanceNorm2d(ch)
    if kind == "batch":
        return nn.BatchNorm2d(ch)
    if kind == "group":
        return nn.GroupNorm(ch // 8, ch)
    return nn.Sequential()


class ResBlock(nn.Module):
    """extractor.py:6-56: two 3x3 convs with a (strided 1x1) skip."""

    def __init__(self, cin: int, cout: int, norm: str, stride: int):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride=stride, padding=1)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.norm1, self.norm2 = _norm(norm, cout), _norm(norm, cout)
        self.downsample = None
        if stride != 1:
            self.norm3 = _norm(norm, cout)
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride=stride), self.norm3)

    def forward(self, x):
        y = F.relu(self.norm1(self.conv1(x)))
        y = F.relu(self.norm2(self.conv2(y)))
        skip = x if self.downsample is None else self.downsample(x)
        return F.relu(skip + y)


class Encoder(nn.Module):
    """extractor.py:118-192 (BasicEncoder): 1/8-resolution feature / context network."""

    def __init__(self, out_dim: int, norm: str):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3)
        self.norm1 = _norm(norm, 64) if norm != "group" else nn.GroupNorm(8, 64)
        widths, strides, cin = (64, 96, 128), (1, 2, 2), 64
        for idx, (w, s) in enumerate(zip(widths, strides), start=1):
            setattr(self, f"layer{idx}", nn.Sequential(ResBlock(cin, w, norm, s), ResBlock(w, w, norm, 1)))
            cin = w
        self.conv2 = nn.Conv2d(128, out_dim, 1)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = F.relu(self.norm1(self.conv1(x)))
        x = self.layer3(self.layer2(self.layer1(x)))
        return self.conv2(x)


class NarrowBlock(nn.Module):
    """extractor.py:58-116 (BottleneckBlock): 1x1 -> 3x3 (strided) -> 1x1 at a quarter of the width, with a
    strided 1x1 skip when the resolution changes."""

    def __init__(self, cin: int, cout: int, norm: str, stride: int):
        super().__init__()
        mid = cout // 4
        self.conv1 = nn.Conv2d(cin, mid, 1)
        self.conv2 = nn.Conv2d(mid, mid, 3, padding=1, stride=stride)
        self.conv3 = nn.Conv2d(mid, cout, 1)
        self.norm1, self.norm2, self.norm3 = _norm(norm, mid), _norm(norm, mid), _norm(norm, cout)
        self.downsample = None
        if stride != 1:
            self.norm4 = _norm(norm, cout)
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride=stride), self.norm4)

    def forward(self, x):
        y = F.relu(self.norm1(self.conv1(x)))
        y = F.relu(self.norm2(self.conv2(y)))
        y = F.relu(self.norm3(self.conv3(y)))
        skip = x if self.downsample is None else self.downsample(x)
        return F.relu(skip + y)


class SmallEncoder(nn.Module):
    """extractor.py:195-267: the 1/8-resolution encoder of the small model (32-64-96 bottleneck stages)."""

    def __init__(self, out_dim: int, norm: str):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 32, 7, stride=2, padding=3)
        self.norm1 = _norm(norm, 32)
        cin = 32
        for idx, (w, stride) in enumerate(((32, 1), (64, 2), (96, 2)), start=1):
            setattr(self, f"layer{idx}", nn.Sequential(NarrowBlock(cin, w, norm, stride), NarrowBlock(w, w, norm, 1)))
            cin = w
        self.conv2 = nn.Conv2d(96, out_dim, 1)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = F.relu(self.norm1(self.conv1(x)))
        return self.conv2(self.layer3(self.layer2(self.layer1(x))))


class SmallMotionEncoder(nn.Module):
    """update.py:62-77: consumes the 196-channel lookup of the small model."""

    def __init__(self, corr_channels: int):
        super().__init__()
        self.convc1 = nn.Conv2d(corr_channels, 96, 1)
        self.convf1 = nn.Conv2d(2, 64, 7, padding=3)
        self.convf2 = nn.Conv2d(64, 32, 3, padding=1)
        self.conv = nn.Conv2d(128, 80, 3, padding=1)

    def forward(self, flow, corr, encoded=False):
        c = corr if encoded else F.relu(self.convc1(corr))
        f = F.relu(self.convf2(F.relu(self.convf1(flow))))
        return torch.cat([F.relu(self.conv(torch.cat([c, f], dim=1))), flow], dim=1)


class GRU3x3(nn.Module):
    """update.py:16-31 (ConvGRU): one 3x3 convolutional GRU step."""

    def __init__(self, hidden: int, inp: int):
        super().__init__()
        for gate in "zrq":
            setattr(self, "conv" + gate, nn.Conv2d(hidden + inp, hidden, 3, padding=1))

    def forward(self, h, x):
        hx = torch.cat([h, x], dim=1)
        z, r = torch.sigmoid(self.convz(hx)), torch.sigmoid(self.convr(hx))
        q = torch.tanh(self.convq(torch.cat([r * h, x], dim=1)))
        return (1 - z) * h + z * q


class SmallUpdateBlock(nn.Module):
    """update.py:99-111: no upsampling mask (the small model upsamples bilinearly, raft.py:134-135)."""

    def __init__(self, corr_channels: int, hidden: int = 96):
        super().__init__()
        self.encoder = SmallMotionEncoder(corr_channels)
        self.gru = GRU3x3(hidden, 82 + 64)
        self.flow_head = FlowHead(hidden, 128)

    def forward(self, net, inp, corr, flow, encoded=False):
        net = self.gru(net, torch.cat([inp, self.encoder(flow, corr, encoded)], dim=1))
        return net, None, self.flow_head(net)


class MotionEncoder(nn.Module):
    """update.py:79-97: consumes the 324-channel lookup (convc1 is a 1x1 conv 324->256)."""

    def __init__(self, corr_channels: int):
        super().__init__()
        self.convc1 = nn.Conv2d(corr_channels, 256, 1)
        self.convc2 = nn.Conv2d(256, 192, 3, padding=1)
        self.convf1 = nn.Conv2d(2, 128, 7, padding=3)
        self.convf2 = nn.Conv2d(128, 64, 3, padding=1)
        self.conv = nn.Conv2d(256, 126, 3, padding=1)

    def forward(self, flow, corr, encoded=False):
        """``encoded``: ``corr`` already is relu(convc1(lookup)) (the fused lookup of the library)."""
        c = F.relu(self.convc2(corr if encoded else F.relu(self.convc1(corr))))
        f = F.relu(self.convf2(F.relu(self.convf1(flow))))
        return torch.cat([F.relu(self.conv(torch.cat([c, f], dim=1))), flow], dim=1)


class SepGRU(nn.Module):
    """update.py:33-60: a 1x5 then a 5x1 convolutional GRU step."""

    def __init__(self, hidden: int, inp: int):
        super().__init__()
        for tag, k, p in (("1", (1, 5), (0, 2)), ("2", (5, 1), (2, 0))):
            for gate in "zrq":
                setattr(self, f"conv{gate}{tag}", nn.Conv2d(hidden + inp, hidden, k, padding=p))

    def _step(self, h, x, tag):
        hx = torch.cat([h, x], dim=1)
        z = torch.sigmoid(getattr(self, "convz" + tag)(hx))
        r = torch.sigmoid(getattr(self, "convr" + tag)(hx))
        q = torch.tanh(getattr(self, "convq" + tag)(torch.cat([r * h, x], dim=1)))
        return (1 - z) * h + z * q

    def forward(self, h, x):
        return self._step(self._step(h, x, "1"), x, "2")


class FlowHead(nn.Module):
    def __init__(self, cin: int, hidden: int):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, hidden, 3, padding=1)
        self.conv2 = nn.Conv2d(hidden, 2, 3, padding=1)

    def forward(self, x):
        return self.conv2(F.relu(self.conv1(x)))


class UpdateBlock(nn.Module):
    """update.py:113-136 (BasicUpdateBlock)."""

    def __init__(self, corr_channels: int, hidden: int = 128):
        super().__init__()
        self.encoder = MotionEncoder(corr_channels)
        self.gru = SepGRU(hidden, 128 + hidden)
        self.flow_head = FlowHead(hidden, 256)
        self.mask = nn.Sequential(nn.Conv2d(128, 256, 3, padding=1), nn.ReLU(inplace=True),
                                  nn.Conv2d(256, 64 * 9, 1))

    def forward(self, net, inp, corr, flow, encoded=False):
        net = self.gru(net, torch.cat([inp, self.encoder(flow, corr, encoded)], dim=1))
        return net, 0.25 * self.mask(net), self.flow_head(net)


def pixel_grid(B: int, H: int, W: int, device) -> torch.Tensor:
    """utils.py:74-77: (B,2,H,W), channel 0 = x, channel 1 = y."""
    ys, xs = torch.meshgrid(torch.arange(H, device=device), torch.arange(W, device=device), indexing="ij")
    return torch.stack([xs, ys], dim=0).float()[None].repeat(B, 1, 1, 1)


def convex_upsample(flow: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """raft.py:72-83: 8x upsampling as a learned convex combination of the 3x3 neighbourhood."""
    B, _, H, W = flow.shape
    w = torch.softmax(mask.view(B, 1, 9, 8, 8, H, W), dim=2)
    nb = F.unfold(8 * flow, [3, 3], padding=1).view(B, 2, 9, 1, 1, H, W)
    up = (w * nb).sum(dim=2)
    return up.permute(0, 1, 4, 2, 5, 3).reshape(B, 2, 8 * H, 8 * W)


def bilinear_upsample8(flow: torch.Tensor) -> torch.Tensor:
    """utils.py:80-82 (upflow8)."""
    return 8 * F.interpolate(flow, size=(8 * flow.shape[2], 8 * flow.shape[3]), mode="bilinear", align_corners=True)


class StockRAFT(nn.Module):
    """raft.py:24-144.  Default: the basic model (hidden 128, context 128, 4 levels, radius 4, D=256);
    ``small=True``: raft.py:29-33,49-51 (hidden 96, context 64, radius 3, D=128, 196 lookup channels, bilinear
    upsampling)."""

    def __init__(self, small: bool = False):
        super().__init__()
        self.small, self.levels = small, 4
        if small:
            self.radius, self.hidden, self.context = 3, 96, 64
            self.fnet = SmallEncoder(128, "instance")
            self.cnet = SmallEncoder(self.hidden + self.context, "none")
            self.update_block = SmallUpdateBlock(self.levels * (2 * self.radius + 1) ** 2, self.hidden)
        else:
            self.radius, self.hidden, self.context = 4, 128, 128
            self.fnet = Encoder(256, "instance")
            self.cnet = Encoder(self.hidden + self.context, "batch")
            self.update_block = UpdateBlock(self.levels * (2 * self.radius + 1) ** 2, self.hidden)

    def features(self, image1, image2):
        """raft.py:89-103 + 110-114: normalise, encode both frames in one batch, context split."""
        image1 = (2 * (image1 / 255.0) - 1.0).contiguous()
        image2 = (2 * (image2 / 255.0) - 1.0).contiguous()
        fmaps = self.fnet(torch.cat([image1, image2], dim=0))
        fmap1, fmap2 = torch.split(fmaps, [image1.shape[0]] * 2, dim=0)
        net, inp = torch.split(self.cnet(image1), [self.hidden, self.context], dim=1)
        return fmap1.float(), fmap2.float(), torch.tanh(net), torch.relu(inp)

    def forward(self, image1, image2, corr_factory, iters: int = 12, flow_init=None, lookup_hook=None,
                upsampler=None, upsampler8=None, convc1_fn=None):
        """Returns (flow at 1/8 resolution, upsampled flow) like raft.py's test_mode.  ``upsampler(flow, mask)``
        replaces raft.py:72-83 (default: the stock-torch op sequence); like the reference it runs after
        every iteration (raft.py:134-139)."""
        upsampler = upsampler or convex_upsample   # ``convc1_fn``: see the loop below
        upsampler8 = upsampler8 or bilinear_upsample8
        fmap1, fmap2, net, inp = self.features(image1, image2)
        corr_fn = corr_factory(fmap1, fmap2, self.radius)
        B, _, H, W = fmap1.shape
        coords0 = pixel_grid(B, H, W, fmap1.device)
        coords1 = coords0.clone() if flow_init is None else coords0 + flow_init
        flow_up = None
        for _ in range(iters):
            coords1 = coords1.detach()
            corr = corr_fn(coords1) if lookup_hook is None else lookup_hook(corr_fn, coords1)
            if convc1_fn is not None:
                # ``convc1_fn(lookup, weight, bias)`` = relu(convc1(lookup)): the library's tensor-core 1x1
                # convolution over a channels-last lookup instead of update.py:83,90
                c1 = self.update_block.encoder.convc1
                corr = convc1_fn(corr, c1.weight, c1.bias)
            net, up_mask, delta = self.update_block(net, inp, corr, coords1 - coords0, convc1_fn is not None)
            coords1 = coords1 + delta
            flow = coords1 - coords0
            flow_up = upsampler8(flow) if up_mask is None else upsampler(flow, up_mask)
        return coords1 - coords0, flow_up
