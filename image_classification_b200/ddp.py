"""Data-parallel gradient exchange: bucketed all-reduce overlapped with backward.

The reference is single-process; this layer is what BASELINE configs 3-4 add.
One process per GPU; the only cross-rank step is the gradient SUM (then / world_size) once per
optimizer step.  Buckets are filled in reverse parameter order (the order backward produces
gradients), each bucket's all-reduce is issued asynchronously the moment its last gradient has
been accumulated — NCCL runs it on its own stream over NVLink 5 / NVSwitch while the remaining
backward kernels keep the SMs busy — and `finish()` joins before unscale/clip/step.  On
non-boundary accumulation micro-steps nothing is sent (the `no_sync` of the reference's
accumulate-2 loop, /root/reference/train_advanced_v4.py:225).  Parameters that never receive a
gradient (the reference's three discarded aux heads, :146-149/:219-220) are left out of the
buckets by `discover_unused`.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class BucketedGradReducer:
    def __init__(self, params, bucket_bytes: int = 32 << 20, process_group=None):
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        self.sync = True
        self._handles = []
        self._hooks = []
        self._build(bucket_bytes)

    # ---- construction -------------------------------------------------------------------
    def _build(self, bucket_bytes: int):
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self.buckets = []          # list of dict(params, flat, views, pending)
        self._where = {}
        cur, cur_bytes = [], 0
        for p in reversed(self.params):
            cur.append(p)
            cur_bytes += p.numel() * p.element_size()
            if cur_bytes >= bucket_bytes:
                self._add_bucket(cur)
                cur, cur_bytes = [], 0
        if cur:
            self._add_bucket(cur)
        for p in self.params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))

    def _add_bucket(self, plist):
        p0 = plist[0]
        n = sum(p.numel() for p in plist)
        flat = torch.zeros(n, device=p0.device, dtype=p0.dtype)
        views, off = [], 0
        for p in plist:
            views.append(flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        b = dict(params=list(plist), flat=flat, views=views, pending=len(plist))
        for p in plist:
            self._where[p] = len(self.buckets)
        self.buckets.append(b)

    def discover_unused(self, bucket_bytes: int = 32 << 20):
        """Call after one backward: parameters whose .grad is still None are dropped from the
        buckets (they would otherwise stall a bucket forever)."""
        used = [p for p in self.params if p.grad is not None]
        if len(used) != len(self.params):
            self.params = used
            self._build(bucket_bytes)

    # ---- per step -----------------------------------------------------------------------
    def set_sync(self, sync: bool):
        self.sync = bool(sync)
        if self.sync:
            for b in self.buckets:
                b["pending"] = len(b["params"])
            self._handles = []

    def _on_grad(self, p):
        if not self.sync or self.world == 1:
            return
        b = self.buckets[self._where[p]]
        b["pending"] -= 1
        if b["pending"] == 0:
            self._launch(b)

    def _launch(self, b):
        torch._foreach_copy_(b["views"], [p.grad for p in b["params"]])
        h = dist.all_reduce(b["flat"], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self._handles.append((h, b))

    def finish(self):
        """Join all in-flight bucket reductions and write the averaged gradients back."""
        if self.world == 1 or not self.sync:
            return
        for b in self.buckets:      # buckets whose hooks did not all fire (defensive)
            if b["pending"] > 0 and all(p.grad is not None for p in b["params"]):
                b["pending"] = 0
                self._launch(b)
        inv = 1.0 / self.world
        for h, b in self._handles:
            h.wait()
            b["flat"].mul_(inv)
            torch._foreach_copy_([p.grad for p in b["params"]], b["views"])
        self._handles = []
