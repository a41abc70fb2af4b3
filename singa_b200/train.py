"""Data-parallel train step around ``SINGA`` (reference train.py:113-133), SURVEY.md 8f-1.

One process per GPU; each rank owns a shard of the complexes (``parallel.shard_by_edges``), runs forward + backward
locally and the only exchange is the gradient all-reduce, launched per bucket from backward
(``parallel.GradientBuckets``).  Differences from the reference loop, each a host-side fix named in SURVEY.md 8f-1:
the batch iterator lives across steps (train.py:116 re-creates the DataLoader iterator, and its worker processes, on
every step); the batch arrives as ONE pinned buffer with its CSRs prebuilt (singa_b200/dataset.py).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.distributed as dist

from . import synthetic as syn
from .dataset import BatchLoader, ComplexDataset, HostBatch
from .parallel import GradientBuckets, shard_by_edges


class Trainer:
    def __init__(self, model, dataset: ComplexDataset, batch_size_per_rank: int, lr: float = 1e-4, betas=(0.99, 0.999),
                 weight_decay: float = 0.0, max_grad_norm: float = math.inf, seed: int = 0, device=None,
                 shuffle: bool = True, prefetch: int = 0, reserve_fraction: float = 0.0):
        self.model = model
        self.device = device or next(model.parameters()).device
        world = dist.get_world_size() if dist.is_initialized() else 1
        rank = dist.get_rank() if dist.is_initialized() else 0
        # equal complex counts per rank: every rank draws batch_size_per_rank per step, so shards of unequal length
        # would wrap around at different steps and a step would not be one global batch
        shard = (shard_by_edges(dataset.edge_counts(), world, equal_counts=True)[rank] if world > 1
                 else list(range(len(dataset))))
        self.loader = BatchLoader(dataset, batch_size_per_rank, indices=shard, seed=seed + rank, shuffle=shuffle,
                                  prefetch=prefetch)
        self.criterion = torch.nn.CrossEntropyLoss()
        self.optimizer = torch.optim.Adam(model.parameters(), lr=lr, betas=betas, weight_decay=weight_decay)
        self.buckets = GradientBuckets(model.named_parameters())
        self.max_grad_norm = max_grad_norm
        self.it = 0
        if reserve_fraction > 0:   # size the allocator pool once: no pool extension (a host stall) inside a step
            from . import ops
            ops.reserve_workspace(reserve_fraction, self.device)

    def step(self, host_batch: Optional[HostBatch] = None):
        """zero_grad -> H2D -> forward -> cross entropy -> backward (+ overlapped all-reduce) -> clip -> Adam
        (train.py:113-127).  Returns (loss, grad_norm) as device tensors (no host sync)."""
        self.model.train()
        self.buckets.zero_grad()
        hb = host_batch if host_batch is not None else next(self.loader)
        db = hb.to(self.device)
        g = db.graph()
        sizes = dict(batch_size=hb.meta["batch_size"], max_atoms=hb.meta["max_p"], max_aa=hb.meta["max_l"])
        logits = self.model(syn.GraphView(g), sizes=sizes, graphs=g.get("enc_graphs"))
        loss = self.criterion(logits, g["smi_tgt"].contiguous().view(-1))
        loss.backward()
        self.buckets.finish()
        grad_norm = torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.max_grad_norm)
        self.optimizer.step()
        self.it += 1
        return loss.detach(), grad_norm
