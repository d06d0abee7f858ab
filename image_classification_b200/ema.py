"""EMA of the parameters as one multi-tensor launch.

Same interface and semantics as the reference's `EMA` class
(/root/reference/train_advanced_v4.py:62-94): `register()` clones every `requires_grad`
parameter into `shadow[name]`, `update()` sets `shadow = (1 - decay) * param + decay * shadow`,
`apply_shadow()` / `restore()` swap `param.data` for evaluation.  The reference's `update()` is a
Python loop of ~3 elementwise kernels + 2 allocations per tensor (~1 400 launches per optimizer
step for ConvNeXt-Base); here it is a single kernel over a device-side pointer table, and the
result is bit-identical (separately rounded multiplies and add, like the eager expression).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


class EMAB200:
    def __init__(self, model: torch.nn.Module, decay: float):
        self.model = model
        self.decay = decay
        self.shadow: dict[str, torch.Tensor] = {}
        self.backup: dict[str, torch.Tensor] = {}
        self._table = None
        self._table_key = None
        self.register()

    # ---- reference API -------------------------------------------------------------------
    def register(self):
        for name, param in self.model.named_parameters():
            if param.requires_grad:
                self.shadow[name] = param.data.clone()
        self._table = None

    def update(self):
        params, shadows = [], []
        for name, param in self.model.named_parameters():
            if param.requires_grad:
                assert name in self.shadow
                params.append(param.data)
                shadows.append(self.shadow[name])
        if not params:
            return
        dev = params[0].device
        if dev.type != "cuda":
            raise _lib.CnxError("EMAB200.update: parameters must live on a CUDA device")
        key = tuple(p.data_ptr() for p in params) + tuple(s.data_ptr() for s in shadows)
        if self._table is None or self._table_key != key:
            self._build_table(params, shadows, dev)
            self._table_key = key
        pp, sp, numels, prefix, T, total = self._table
        # the reference evaluates (1.0 - decay) in float64, then the product in float32
        omd = float(np.float32(1.0 - self.decay))
        d = float(np.float32(self.decay))
        _lib.call("cnx_ema_lerp_multi", pp.data_ptr(), sp.data_ptr(), numels.data_ptr(),
                  prefix.data_ptr(), T, total, omd, d, torch.cuda.current_stream(dev).cuda_stream)

    def apply_shadow(self):
        for name, param in self.model.named_parameters():
            if param.requires_grad:
                assert name in self.shadow
                self.backup[name] = param.data
                param.data = self.shadow[name]

    def restore(self):
        for name, param in self.model.named_parameters():
            if param.requires_grad:
                assert name in self.backup
                param.data = self.backup[name]
        self.backup = {}

    # ---- device table --------------------------------------------------------------------
    def _build_table(self, params, shadows, dev):
        chunk = _lib.load().cnx_ema_chunk_elems()
        for p, s in zip(params, shadows):
            if p.dtype != torch.float32 or s.dtype != torch.float32:
                raise _lib.CnxError("EMAB200: parameters and shadows must be float32")
            # the lerp is elementwise over the storage: any dense layout works (channels_last
            # conv weights included) as long as parameter and shadow share it
            dense = p.is_contiguous() or (p.dim() == 4 and
                                          p.is_contiguous(memory_format=torch.channels_last))
            if not dense or p.stride() != s.stride():
                raise _lib.CnxError("EMAB200: parameters must be dense and laid out like their "
                                    f"shadows (got strides {p.stride()} vs {s.stride()})")
        numel = np.array([p.numel() for p in params], dtype=np.int64)
        nchunks = (numel + chunk - 1) // chunk
        prefix = np.zeros(len(params) + 1, dtype=np.int64)
        np.cumsum(nchunks, out=prefix[1:])
        pp = np.array([p.data_ptr() for p in params], dtype=np.int64)
        sp = np.array([s.data_ptr() for s in shadows], dtype=np.int64)
        to_dev = lambda a: torch.from_numpy(a).to(dev)
        self._table = (to_dev(pp), to_dev(sp), to_dev(numel), to_dev(prefix), len(params),
                       int(prefix[-1]))


class AveragedModelB200(torch.nn.Module):
    """`torch.optim.swa_utils.AveragedModel(model)` with its default equal-weight average, as the
    older reference loop uses it for SWA (/root/reference/previous/V3-efb2/train_advanced_v3.0.py:445
    construction, :505-507 `swa_model.update_parameters(model)` once per epoch after `swa_start`):
    same attributes (`module`, `n_averaged`), same state_dict keys, same result
    `avg <- avg + (p - avg) / (n + 1)`  (the first call copies), but ONE `cnx_ema_lerp_multi` launch
    over all parameters instead of a foreach pass per call.  Buffers (BatchNorm statistics) are not
    averaged, as with torch's default `use_buffers=False`."""

    def __init__(self, model: torch.nn.Module):
        super().__init__()
        import copy
        self.module = copy.deepcopy(model)
        dev = next(model.parameters()).device
        self.register_buffer("n_averaged", torch.tensor(0, dtype=torch.long, device=dev))
        self._n = 0                        # host mirror of n_averaged: no device read per update
        self._lerp = EMAB200.__new__(EMAB200)
        self._lerp._table, self._lerp._table_key = None, None

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def update_parameters(self, model: torch.nn.Module) -> None:
        params = [p.data for p in model.parameters()]
        avgs = [p.data for p in self.module.parameters()]
        if len(params) != len(avgs):
            raise _lib.CnxError("AveragedModelB200: the model does not match the averaged copy")
        dev = params[0].device
        if dev.type != "cuda":
            raise _lib.CnxError("AveragedModelB200.update_parameters: parameters must live on a "
                                "CUDA device (no CPU path)")
        key = tuple(p.data_ptr() for p in params) + tuple(s.data_ptr() for s in avgs)
        if self._lerp._table is None or self._lerp._table_key != key:
            self._lerp._build_table(params, avgs, dev)
            self._lerp._table_key = key
        pp, sp, numels, prefix, T, total = self._lerp._table
        n = self._n
        w = 1.0 if n == 0 else 1.0 / (n + 1)               # weight of the new parameters
        _lib.call("cnx_ema_lerp_multi", pp.data_ptr(), sp.data_ptr(), numels.data_ptr(),
                  prefix.data_ptr(), T, total, float(np.float32(w)), float(np.float32(1.0 - w)),
                  torch.cuda.current_stream(dev).cuda_stream)
        self._n += 1
        self.n_averaged += 1
