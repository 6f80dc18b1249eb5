"""BASELINE.json configs[4] as written: a whole RAFT training step — forward (12 refinement iterations, list of
12 upsampled predictions), ``sequence_loss``, backward, gradient clipping, AdamW + OneCycle step — at 368x496 with
batch 12 per GPU, one process per GPU, gradients all-reduced by ``DistributedDataParallel`` over NCCL (the
reference wraps the model in ``nn.DataParallel``, train.py:138; the all-reduce is outside the optimised path).

The model is the UNMODIFIED reference ``core/raft.py`` (baseline/_ref, tools/install_ref.py) and the loss / optimiser
recipe are the reference's own ``sequence_loss`` / ``fetch_optimizer`` (train.py:47-86), with the hyper-parameters of
its FlyingChairs stage (train_standard.sh: lr 4e-4, wdecay 1e-4, clip 1.0, gamma 0.8).  ``dropin=True`` runs it with
this repository's ``corr.py`` ahead of ``core/`` on ``sys.path`` (CorrBlock forward + backward in
libraftcorr_b200.so), ``dropin=False`` with the reference's own ``corr.py`` (ATen ops).  Used by bench.py and the
tests; nothing here is on the product path."""
from __future__ import annotations

import argparse

import torch

from harness import ref_model

IMG_H, IMG_W, ITERS, BATCH_PER_GPU = 368, 496, 12, 12


def chairs_args(**model_flags) -> argparse.Namespace:
    a = ref_model.make_args(**model_flags)
    a.lr, a.wdecay, a.epsilon, a.num_steps, a.clip, a.gamma, a.iters = 4e-4, 1e-4, 1e-8, 100000, 1.0, 0.8, ITERS
    return a


def synthetic_batch(batch: int, device, seed: int, size=(IMG_H, IMG_W)):
    gen = torch.Generator().manual_seed(seed)
    h, w = size
    im1 = torch.rand(batch, 3, h, w, generator=gen) * 255.0
    im2 = torch.rand(batch, 3, h, w, generator=gen) * 255.0
    flow = 5.0 * torch.randn(batch, 2, h, w, generator=gen)
    valid = torch.ones(batch, h, w)
    return tuple(t.to(device) for t in (im1, im2, flow, valid))


class Trainer:
    def __init__(self, device, dropin: bool, distributed: bool = False, local_rank: int = 0, seed: int = 1234,
                 precision: str | None = None, iters: int = ITERS):
        import os
        if precision is not None:
            os.environ["RAFTCORR_PRECISION"] = precision
        self.raft = ref_model.load_raft_module(dropin=dropin)
        self.sequence_loss, fetch_optimizer = ref_model.load_train_functions()
        self.args = chairs_args()
        self.args.iters = iters
        torch.manual_seed(seed)
        self.net = self.raft.RAFT(self.args).to(device).train()      # chairs stage: batch norm is not frozen
        self.model = self.net
        if distributed:
            from torch.nn.parallel import DistributedDataParallel
            on_gpu = torch.device(device).type == "cuda"
            self.model = DistributedDataParallel(self.net, device_ids=[local_rank] if on_gpu else None)
        self.optimizer, self.scheduler = fetch_optimizer(self.args, self.model)

    def step(self, im1, im2, flow, valid):
        """train.py:163-181 without the data loader, the noise augmentation and the logger."""
        self.optimizer.zero_grad()
        preds = self.model(im1, im2, iters=self.args.iters)
        loss, metrics = self.sequence_loss(preds, flow, valid, self.args.gamma)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.args.clip)
        self.optimizer.step()
        self.scheduler.step()
        return loss, metrics
