"""Data-parallel gradient exchange: bucketed all-reduce overlapped with backward.

The reference is single-process; this layer is what BASELINE configs 3-4 add.
One process per GPU; the only cross-rank step is the gradient SUM (then / world_size) once per
optimizer step.  Buckets are filled in reverse parameter order (the order backward produces
gradients), each bucket's all-reduce is issued asynchronously the moment its last gradient has
been accumulated — NCCL runs it on its own stream over NVLink 5 / NVSwitch while the remaining
backward kernels keep the SMs busy — and `finish()` joins before unscale/clip/step.  On
non-boundary accumulation micro-steps nothing is sent (the `no_sync` of the reference's
accumulate-2 loop, /root/reference/train_advanced_v4.py:225).

Parameters that never receive a gradient (the reference's three discarded aux heads,
:146-149/:219-220) are found on the FIRST synchronised backward: `finish()` reduces an incomplete
bucket with zeros in the place of the missing gradients (every rank runs the same model, so every
rank sees the same set) and then rebuilds the buckets without those parameters.  A parameter that
was dropped this way and later does receive a gradient on a synchronised step raises — ranks must
never diverge silently.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class BucketedGradReducer:
    def __init__(self, params, bucket_bytes: int = 32 << 20, process_group=None):
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("BucketedGradReducer: no parameter requires a gradient")
        self.bucket_bytes = bucket_bytes
        self.sync = True
        self.discovered = False        # set once the unused-parameter set is known
        self.dropped = []              # parameters found unused (no gradient on a sync backward)
        self.reduced_steps = 0         # synchronised steps joined by finish() (bench asserts > 0)
        self._handles = []
        self._hooks = []
        self._build()

    # ---- construction -------------------------------------------------------------------
    def _build(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self.buckets = []          # list of dict(params, flat, views, pending)
        self._where = {}
        cur, cur_bytes = [], 0
        for p in reversed(self.params):
            cur.append(p)
            cur_bytes += p.numel() * p.element_size()
            if cur_bytes >= self.bucket_bytes:
                self._add_bucket(cur)
                cur, cur_bytes = [], 0
        if cur:
            self._add_bucket(cur)
        for p in self.params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        for p in self.dropped:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_dropped_grad))

    def _add_bucket(self, plist):
        p0 = plist[0]
        n = sum(p.numel() for p in plist)
        flat = torch.zeros(n, device=p0.device, dtype=p0.dtype)
        views, off = [], 0
        for p in plist:
            # a view with the parameter's own (dense) memory order: channels_last conv weights keep
            # gradients laid out like the parameter, which the fused optimizer step requires
            dense = p.is_contiguous() or (p.dim() == 4 and
                                          p.is_contiguous(memory_format=torch.channels_last))
            chunk = flat[off:off + p.numel()]
            views.append(chunk.as_strided(p.size(), p.stride()) if dense else chunk.view_as(p))
            off += p.numel()
        b = dict(params=list(plist), flat=flat, views=views, pending=len(plist))
        for p in plist:
            self._where[p] = len(self.buckets)
        self.buckets.append(b)

    def discover_unused(self):
        """Explicit form of what the first synchronised `finish()` does on its own: call right
        after a backward that produced every gradient of an optimizer step (i.e. on the boundary
        micro-batch, BEFORE `optimizer.zero_grad()`); parameters whose `.grad` is None are dropped
        from the buckets.  Raises if that would drop every parameter (a call after zero_grad)."""
        used = [p for p in self.params if p.grad is not None]
        if not used:
            raise RuntimeError("BucketedGradReducer.discover_unused: no parameter holds a gradient "
                               "— call it after backward of the boundary micro-batch and before "
                               "optimizer.zero_grad()")
        self._drop([p for p in self.params if p.grad is None])
        self.discovered = True

    def _drop(self, unused):
        if unused:
            gone = set(unused)
            self.dropped += unused
            self.params = [p for p in self.params if p not in gone]
            self._build()

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

    def _on_dropped_grad(self, p):
        if self.sync and self.world > 1:
            raise RuntimeError("BucketedGradReducer: a parameter that had no gradient when the "
                               "buckets were built received one on a synchronised step; it would "
                               "not be all-reduced (rebuild the reducer)")

    def _launch(self, b):
        src = [(v, p.grad) for v, p in zip(b["views"], b["params"]) if p.grad.data_ptr() != v.data_ptr()]
        if src:
            torch._foreach_copy_([v for v, _ in src], [g for _, g in src])
        h = dist.all_reduce(b["flat"], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self._handles.append((h, b))

    def _launch_partial(self, b):
        """A bucket some of whose parameters hold no gradient: zeros stand in for them, so the
        collective keeps the same size on every rank."""
        b["flat"].zero_()
        have = [(v, p.grad) for v, p in zip(b["views"], b["params"])
                if p.grad is not None and p.grad.data_ptr() != v.data_ptr()]
        if have:
            torch._foreach_copy_([v for v, _ in have], [g for _, g in have])
        h = dist.all_reduce(b["flat"], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self._handles.append((h, b))

    def finish(self):
        """Join all in-flight bucket reductions and write the averaged gradients back."""
        if self.world == 1 or not self.sync:
            return
        unused = []
        for b in self.buckets:
            if b["pending"] > 0:
                missing = [p for p in b["params"] if p.grad is None]
                if missing and self.discovered:
                    raise RuntimeError(
                        f"BucketedGradReducer.finish: {len(missing)} parameter(s) of a bucket have "
                        "no gradient on a synchronised step although they had one when the buckets "
                        "were built; the step would leave ranks out of sync")
                unused += missing
                b["pending"] = 0
                self._launch_partial(b) if missing else self._launch(b)
        inv = 1.0 / self.world
        for h, b in self._handles:
            h.wait()
            b["flat"].mul_(inv)
            # no copy back: the parameter's .grad becomes the averaged slice of the bucket (the
            # next backward allocates fresh gradients after zero_grad(set_to_none=True), or
            # accumulates into the slice, which the next step's copy-in then finds in place)
            for v, p in zip(b["views"], b["params"]):
                if p.grad is not None:
                    p.grad = v
        self._handles = []
        self.reduced_steps += 1
        if not self.discovered:
            # first synchronised backward: the parameters without a gradient are the unused ones
            self.discovered = True
            self._drop(unused)
