"""Synthetic-data, data-parallel training step around the spectral mixers.

Follows the reference's single-GPU loop (``spectre_vit/repl/train.py``: model build :48-59, AdamW
:199-201, autocast forward :219-220, CE outside autocast :226, ``zero_grad(set_to_none=True)`` /
backward / step :235-238; KD branch :298-353) with the B200 changes SURVEY.md section 5 lists:
bf16 autocast instead of fp16 + GradScaler, no per-step ``loss.item()`` host sync, one process per
GPU with an NCCL gradient all-reduce (DistributedDataParallel buckets overlapped with backward --
the only collective on the path; the reference has none).
"""
from __future__ import annotations

import os
from types import SimpleNamespace

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from .model import SpectreViT


def build_model(cfg: SimpleNamespace, mixer: str | None = None) -> SpectreViT:
    """``SpectreViT`` from a reference-format config (train.py:48-59) plus the ``mixer`` key."""
    return SpectreViT(
        img_size=cfg.img_size, patch_size=cfg.patch_size, in_channels=cfg.in_channels, num_classes=cfg.num_classes,
        embed_dim=cfg.embed_dim, num_encoders=cfg.num_encoders, num_heads=cfg.num_heads, hidden_dim=cfg.hidden_dim,
        dropout=cfg.dropout, activation=cfg.activation, mixer=mixer or getattr(cfg, "mixer", "permut"),
    )


class TeacherViT(nn.Module):
    """Random-init, frozen attention ViT used as the distillation teacher of config 5.  The reference
    loads a DINOv3 backbone + linear head (``spectre_vit/distillation.py:5-22``); that third-party
    submodule does not travel, so this is a stock ViT of the same sizes (library SDPA / cuBLAS)."""

    def __init__(self, img_size=224, patch_size=16, in_channels=3, embed_dim=768, depth=12, heads=12, num_classes=1000):
        super().__init__()
        n = (img_size // patch_size) ** 2
        self.patch = nn.Conv2d(in_channels, embed_dim, patch_size, patch_size)
        self.cls = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos = nn.Parameter(0.02 * torch.randn(1, n + 1, embed_dim))
        layer = nn.TransformerEncoderLayer(embed_dim, heads, 4 * embed_dim, dropout=0.0, activation="gelu",
                                           batch_first=True, norm_first=True)
        self.blocks = nn.TransformerEncoder(layer, depth, enable_nested_tensor=False)
        self.norm = nn.LayerNorm(embed_dim)
        self.linear = nn.Linear(embed_dim, num_classes)

    def forward(self, x, return_features=False):
        t = self.patch(x).flatten(2).transpose(1, 2)
        t = torch.cat([self.cls.expand(t.shape[0], -1, -1).to(t.dtype), t], 1) + self.pos.to(t.dtype)
        feat = self.norm(self.blocks(t))[:, 0]
        logits = self.linear(feat)
        return (logits, feat) if return_features else logits


def kd_loss(student_logits, teacher_logits, labels, T=2.0, soft_w=0.25, ce_w=0.75):
    """``0.25 * T^2 * sum(p_T (log p_T - log_softmax(s/T))) / B + 0.75 * CE`` (train.py:334-348)."""
    soft_t = F.softmax(teacher_logits.float() / T, dim=-1)
    log_s = F.log_softmax(student_logits.float() / T, dim=-1)
    soft = torch.sum(soft_t * (soft_t.clamp_min(1e-30).log() - log_s)) / log_s.shape[0] * (T * T)
    return soft_w * soft + ce_w * F.cross_entropy(student_logits.float(), labels)


def dist_env():
    """(rank, local_rank, world) from torchrun's environment; (0, 0, 1) when launched plainly."""
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def init_distributed(backend: str | None = None):
    rank, local_rank, world = dist_env()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            dist.init_process_group(backend, device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group(backend)
    return rank, local_rank, world


class Trainer:
    """One optimisation step = forward (autocast when ``precision == "bf16"``) -> loss -> backward
    (DDP all-reduce overlapped) -> fused AdamW.  ``step`` returns the loss tensor without syncing."""

    def __init__(self, cfg: SimpleNamespace, device, *, mixer: str | None = None, model: nn.Module | None = None,
                 ddp: bool | None = None, seed: int | None = None, teacher: nn.Module | None = None,
                 cuda_graph: bool = False):
        self.cfg, self.device = cfg, torch.device(device)
        torch.manual_seed(cfg.random_seed if seed is None else seed)
        self.model = (model if model is not None else build_model(cfg, mixer)).to(self.device)
        self.bf16 = getattr(cfg, "precision", "fp32") == "bf16" and self.device.type == "cuda"
        self.teacher = None
        if getattr(cfg, "distill", False):
            # `teacher`: any frozen module with forward(img) -> logits (or (logits, features) with return_features=True),
            # e.g. the reference's DinoClassifier(dinov3_vitb16(pretrained=False)) (spectre_vit/distillation.py:5-22,
            # repl/train.py:60-83) built by the caller; default: a stock ViT of the same sizes
            if teacher is None:
                teacher = TeacherViT(cfg.img_size, cfg.patch_size, cfg.in_channels, cfg.teacher_embed_dim,
                                     cfg.teacher_depth, cfg.teacher_heads, cfg.num_classes)
            self.teacher = teacher.to(self.device).eval()
            for p in self.teacher.parameters():
                p.requires_grad_(False)
        world = dist.get_world_size() if dist.is_initialized() else 1
        self.world = world
        self.net = self.model
        # cuda_graph: the whole step (forward, loss, backward, gradient all-reduce, AdamW) is captured once and
        # replayed -- the C ABI only enqueues work on the current stream, so it is capturable as it stands.  The small
        # configs (C1-C3: ~180 launches in a few ms) are launch-bound in eager mode.  Gradients are all-reduced as one
        # flat NCCL call inside the graph instead of DDP buckets.
        self.cuda_graph = bool(cuda_graph) and self.device.type == "cuda"
        self._graph = None
        if (ddp if ddp is not None else world > 1) and world > 1 and not self.cuda_graph:
            ids = [self.device.index] if self.device.type == "cuda" else None
            self.net = nn.parallel.DistributedDataParallel(self.model, device_ids=ids, gradient_as_bucket_view=True,
                                                           static_graph=True)
        self.opt = torch.optim.AdamW(self.model.parameters(), lr=cfg.learning_rate, betas=cfg.adam_betas,
                                     weight_decay=cfg.adam_weight_decay, fused=self.device.type == "cuda",
                                     capturable=self.cuda_graph)

    def synthetic_batch(self, batch: int, rank: int = 0, pinned_host: bool = False):
        """Synthetic data of the configured shape: ``randn`` images, uniform labels, seed 42 + rank."""
        g = torch.Generator().manual_seed(self.cfg.random_seed + rank)
        img = torch.randn(batch, self.cfg.in_channels, self.cfg.img_size, self.cfg.img_size, generator=g)
        lab = torch.randint(0, self.cfg.num_classes, (batch,), generator=g)
        if pinned_host:
            return img.pin_memory(), lab.pin_memory()
        return img.to(self.device), lab.to(self.device)

    def forward_loss(self, img, lab):
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
            if self.teacher is None:
                logits = self.net(img)
            else:
                logits, _ = self.net(img, return_features=True)
                with torch.no_grad():
                    t_logits = self.teacher(img)
        if self.teacher is None:
            return F.cross_entropy(logits.float(), lab)
        c = self.cfg
        return kd_loss(logits, t_logits, lab, c.distill_temperature, c.distill_soft_weight, c.distill_ce_weight)

    def _allreduce_grads(self):
        """One flat all-reduce (mean) of every gradient: the graph-mode stand-in for DDP's buckets."""
        grads = [p.grad for p in self.model.parameters() if p.grad is not None]
        flat = torch.cat([g.reshape(-1).float() for g in grads])
        if dist.get_backend() == "nccl":
            dist.all_reduce(flat, op=dist.ReduceOp.AVG)
        else:                                   # gloo (CPU tests) has no AVG
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
            flat /= dist.get_world_size()
        torch._foreach_copy_(grads, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in grads]), grads)])

    def _eager_step(self, img, lab):
        loss = self.forward_loss(img, lab)
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        if self.cuda_graph and self.world > 1:
            self._allreduce_grads()
        self.opt.step()
        return loss.detach()

    def _capture(self, img, lab):
        from . import _lib

        self._g_img, self._g_lab = torch.empty_like(img), torch.empty_like(lab)
        self._g_img.copy_(img)
        self._g_lab.copy_(lab)
        self._seed_word = torch.zeros(1, dtype=torch.int64, device=self.device)
        _lib.graph_seed["tensor"] = self._seed_word            # dropout seeds: host part frozen, this word bumped per replay
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):                          # warm-up off the capture: lazy inits, allocator, autotune
            for _ in range(3):
                self._seed_word += 1
                self._eager_step(self._g_img, self._g_lab)
        torch.cuda.current_stream(self.device).wait_stream(side)
        before = _lib.launch_counter["kernels"]
        self._graph = torch.cuda.CUDAGraph()
        self.opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(self._graph):
            self._seed_word += 1
            self._g_loss = self._eager_step(self._g_img, self._g_lab)
        self._launches_per_replay = _lib.launch_counter["kernels"] - before

    def step(self, img, lab):
        self.net.train()
        if not self.cuda_graph:
            return self._eager_step(img, lab)
        from . import _lib

        if self._graph is None:
            self._capture(img, lab)
        if img.data_ptr() != self._g_img.data_ptr():
            self._g_img.copy_(img, non_blocking=True)
            self._g_lab.copy_(lab, non_blocking=True)
        self._graph.replay()
        _lib.launch_counter["kernels"] += self._launches_per_replay
        return self._g_loss


class HostBatchPrefetcher:
    """Double-buffered pinned-host -> device feed for ``Trainer.step`` (what a DataLoader with ``pin_memory`` plus
    ``non_blocking`` copies amounts to, reference: ``spectre_vit/repl/train.py:147-173,214-216``): the copy of batch
    i+1 runs on a side stream while step i computes, so every step still moves one whole batch over PCIe but the
    transfer is off the critical path.  ``next()`` returns device tensors valid until the call after next."""

    def __init__(self, device, like_img: torch.Tensor, like_lab: torch.Tensor):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)
        self.bufs = [(torch.empty_like(like_img, device=self.device), torch.empty_like(like_lab, device=self.device))
                     for _ in range(2)]
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        self.i = 0          # index of the next batch handed out
        self.staged = 0     # number of batches whose copy has been enqueued
        for e in self.consumed:
            e.record(torch.cuda.current_stream(self.device))

    def stage(self, himg: torch.Tensor, hlab: torch.Tensor) -> None:
        """Enqueue the H2D copy of one (pinned) host batch on the side stream."""
        b = self.staged & 1
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.consumed[b])
            self.bufs[b][0].copy_(himg, non_blocking=True)
            self.bufs[b][1].copy_(hlab, non_blocking=True)
            self.ready[b].record(self.stream)
        self.staged += 1

    def next(self):
        """Device tensors of the oldest staged batch (the compute stream waits for its copy)."""
        if self.i >= self.staged:
            raise RuntimeError("HostBatchPrefetcher.next(): nothing staged")
        b = self.i & 1
        torch.cuda.current_stream(self.device).wait_event(self.ready[b])
        self.i += 1
        return self.bufs[b]

    def release(self) -> None:
        """Mark the batch returned by the last ``next()`` as consumed (call after the step was enqueued)."""
        self.consumed[(self.i - 1) & 1].record(torch.cuda.current_stream(self.device))

