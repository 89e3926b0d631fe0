"""D1 — data-parallel gradient averaging, overlapped bucket by bucket with backward.

The README's motivation for all_reduce is "send all gradients to all GPUs, average them,
and then perform a local optimizer step" (README.md:5) and for broadcast "share the
parameters" (README.md:286); the reference ships no code for either. Here:

  * `broadcast_parameters` copies rank 0's parameters to every rank (K4);
  * `BucketedGradAverager` lays every parameter's `.grad` out as a view into a few large
    flat buckets, and as soon as backward has produced the last gradient of a bucket it
    enqueues ONE fused all_reduce x 1/world (K2) for that bucket on a high-priority side
    stream, so communication of layer i overlaps the backward GEMMs of layers < i.
    `wait()` makes the compute stream wait for the averages before the optimizer step.

No host synchronisation anywhere: hooks only record/wait CUDA events.
"""
from __future__ import annotations

import os
from typing import Iterable, List, Optional

import torch

from .. import distributed as dist


def broadcast_parameters(module: torch.nn.Module, src: int = 0, group=None) -> None:
    """Every rank ends with global rank `src`'s parameters and buffers."""
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=group)


class _Bucket:
    def __init__(self, params: List[torch.nn.Parameter], dtype: torch.dtype, device, group=None,
                 symmetric: bool = True):
        self.params = params
        self.numel = sum(p.numel() for p in params)
        # Gradient buckets live in symmetric memory when the native backend is up: the fused
        # average then runs as multimem.ld_reduce / multimem.st straight on the bucket
        # (csrc/sym.cuh), no staging copy on either side. Padded to whole 16-byte units.
        pad = 16 // torch.empty(0, dtype=dtype).element_size()
        padded = (self.numel + pad - 1) // pad * pad
        if (symmetric and device.type == "cuda" and dist.is_initialized()
                and dist.get_backend() == "b200" and dist.get_world_size(group) > 1):
            self.flat = dist.empty_symmetric(padded, dtype=dtype, group=group)
            self.flat.zero_()
            self.symmetric = dist.is_symmetric(self.flat, group)
        else:
            self.flat = torch.zeros(padded, dtype=dtype, device=device)
            self.symmetric = False
        self.pending = len(params)
        self.ready = torch.cuda.Event() if device.type == "cuda" else None
        self.done = torch.cuda.Event() if device.type == "cuda" else None
        off = 0
        for p in params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()


class BucketedGradAverager:
    """Gradient buckets + overlapped fused averaging for one module.

    bucket_bytes: target bucket size; the default keeps one 8192x8192 bf16 layer
    (128 MiB) per bucket, which is large enough to run the two-shot/NVLS kernels at full
    bandwidth and small enough that only the last bucket's reduction is exposed.
    """

    def __init__(self, params: Iterable[torch.nn.Parameter], group=None,
                 bucket_bytes: int = 128 << 20, overlap: bool = True, symmetric: bool = True,
                 comm_ctas: int = 0):
        params = [p for p in params if p.requires_grad]
        if not params:
            raise ValueError("no trainable parameters")
        self.group = group
        self.overlap = overlap
        # CTAs of each averaging kernel. The kernels keep one CTA per SM resident for their
        # whole run; while backward is still producing gradients the GEMMs should keep most
        # of the SMs, so overlapped averages run on a handful (measured on 4 GPUs: exposed communication 0.53 ms
        # with 8 CTAs, 0.69 with 32, 1.27 with 148; DDP on NCCL 1.01).
        self.comm_ctas = comm_ctas if comm_ctas > 0 else (int(os.environ.get("PDCC_DP_CTAS", "8")) if overlap else 0)
        self.device = params[0].device
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.buckets: List[_Bucket] = []
        # reverse order: backward produces the last layer's gradients first
        cur: List[torch.nn.Parameter] = []
        cur_bytes = 0
        for p in reversed(params):
            nbytes = p.numel() * p.element_size()
            if cur and (cur_bytes + nbytes > bucket_bytes or p.dtype != cur[0].dtype):
                self.buckets.append(_Bucket(cur, cur[0].dtype, self.device, group, symmetric))
                cur, cur_bytes = [], 0
            cur.append(p)
            cur_bytes += nbytes
        if cur:
            self.buckets.append(_Bucket(cur, cur[0].dtype, self.device, group, symmetric))
        self._stream: Optional[torch.cuda.Stream] = None
        if self.device.type == "cuda" and overlap:
            self._stream = torch.cuda.Stream(device=self.device, priority=-1)
        self._handles = []
        for b in self.buckets:
            for p in b.params:
                self._handles.append(p.register_post_accumulate_grad_hook(self._make_hook(b)))

    def _make_hook(self, b: _Bucket):
        def hook(_param):
            b.pending -= 1
            if b.pending == 0:
                self._launch(b)
        return hook

    def _ctas(self, b: _Bucket) -> int:
        # the CTA cap is for the zero-copy kernels (the switch does the work, a handful of
        # CTAs keep it fed); the staged kernels need their SMs (8 GPUs: 10.9 ms/step with 8)
        return self.comm_ctas if b.symmetric else 0

    def _launch(self, b: _Bucket) -> None:
        if self.world == 1:
            return
        if self._stream is None:
            dist.all_reduce_avg(b.flat, group=self.group, nblocks=self._ctas(b),
                                algo="sym" if b.symmetric else None)
            return
        b.ready.record()  # on the compute stream: this bucket's gradients are complete
        self._stream.wait_event(b.ready)
        with torch.cuda.stream(self._stream):
            dist.all_reduce_avg(b.flat, group=self.group, nblocks=self._ctas(b),
                                algo="sym" if b.symmetric else None)
            b.done.record()

    def zero_grad(self) -> None:
        """Call before every forward: clears the buckets and re-arms the hooks. The grads
        stay views into the buckets (never set them to None)."""
        for b in self.buckets:
            b.flat.zero_()
            b.pending = len(b.params)

    def wait(self) -> None:
        """Compute stream waits for every bucket's average; call before optimizer.step()."""
        if self._stream is None or self.world == 1:
            return
        cur = torch.cuda.current_stream()
        for b in self.buckets:
            cur.wait_event(b.done)

    def remove(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles = []
