"""D1 — the data-parallel MLP of BASELINE.json config 5: 4 linear layers, hidden 8192,
bf16, gradients averaged across ranks by the fused all_reduce kernel while backward runs.

The GEMMs are torch/cuBLAS (collectives are the product here); what this
library adds around them is the communication: parameter broadcast, bucketed fused
averaging on a side stream, and the event handshake before the optimizer step.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import distributed as dist
from ..parallel.data_parallel import BucketedGradAverager, broadcast_parameters


class MLP(nn.Module):
    def __init__(self, hidden: int = 8192, layers: int = 4, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.layers = nn.ModuleList(
            nn.Linear(hidden, hidden, bias=True, dtype=dtype, device=device) for _ in range(layers))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for i, lin in enumerate(self.layers):
            x = lin(x)
            if i + 1 < len(self.layers):
                x = torch.relu(x)
        return x


class DataParallelTrainer:
    """One training step = zero buckets, forward, backward (bucket averages start as soon as
    each bucket's gradients exist), wait, local SGD step."""

    def __init__(self, model: nn.Module, lr: float = 1e-3, group=None, overlap: bool = True,
                 bucket_bytes: int = 128 << 20, sync_params: bool = True, symmetric: bool = True,
                 comm_ctas: int = 0):
        self.model = model
        if sync_params and dist.is_initialized() and dist.get_world_size(group) > 1:
            broadcast_parameters(model, src=dist.get_global_rank(group, 0) if group else 0,
                                 group=group)
        self.avg = BucketedGradAverager(model.parameters(), group=group, overlap=overlap,
                                        bucket_bytes=bucket_bytes, symmetric=symmetric,
                                        comm_ctas=comm_ctas)
        self.opt = torch.optim.SGD(model.parameters(), lr=lr)

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        self.avg.zero_grad()
        out = self.model(x)
        loss = torch.nn.functional.mse_loss(out.float(), y.float())
        loss.backward()
        self.avg.wait()
        self.opt.step()
        return loss.detach()


def smoke_step(hidden: int = 256, layers: int = 2, batch: int = 64) -> float:
    """Tiny forward + backward + gradient post-processing + SGD on the current GPU, without a
    process group: the world-size-1 form of the DP step (the fused average degenerates to
    the local scale+cast kernel)."""
    from ..ops import collectives as K

    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(0)
    model = MLP(hidden, layers, device=dev)
    tr = DataParallelTrainer(model, sync_params=False)
    x = torch.randn(batch, hidden, device=dev, dtype=torch.bfloat16)
    y = torch.randn(batch, hidden, device=dev, dtype=torch.bfloat16)
    l0 = float(tr.step(x, y))
    master = torch.empty(tr.avg.buckets[0].numel, device=dev)
    K.local_scale_cast(master, tr.avg.buckets[0].flat, 1.0)  # K2 at world size 1
    assert torch.isfinite(master).all()
    l1 = float(tr.step(x, y))
    assert l1 == l1 and l0 == l0
    return l1
