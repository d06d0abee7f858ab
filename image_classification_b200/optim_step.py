"""The optimizer step of the reference loop in two launches.

Replaces, on the boundary micro-batch of `train_one_epoch`
(/root/reference/train_advanced_v4.py:226-237):

    scaler.unscale_(optimizer)                                   # :226
    torch.nn.utils.clip_grad_norm_(model.parameters(), clip)     # :228
    scaler.step(optimizer)                                       # :229  (AdamW, :504-510)
    scaler.update()                                              # :230
    ...
    ema.update()                                                 # :236-237  (EMA.update :75-80)

with `FusedOptimStep(model, optimizer, scaler, ema, clip).step()`:
`cnx_optim_grad_norm` reads every gradient once (inf/nan check of the unscaled values + global
norm), `cnx_optim_adamw_ema` applies `grad * clip_coef / scale` on the fly inside the AdamW update
and lerps the EMA shadow with the new parameter in the same pass; the loss scale is then updated
with torch's own `_amp_update_scale_` from the device-side found-inf flag — no host sync anywhere.
The `torch.optim.AdamW` object stays the owner of the hyper-parameters and of the state
(`exp_avg`, `exp_avg_sq`, `step`), so LR schedulers and `optimizer.state_dict()` keep working.
`optimizer.zero_grad()` and `scheduler.step()` remain the caller's (they launch nothing).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .block import notify_weights_changed


def _same_layout(a: torch.Tensor, b: torch.Tensor) -> bool:
    """Same memory order: equal strides on every dimension that has more than one element."""
    return a.shape == b.shape and all(sa == sb for n, sa, sb in zip(a.shape, a.stride(), b.stride())
                                      if n > 1)


def _dense(t: torch.Tensor) -> bool:
    return t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last))


class _Table:
    """Device pointer tables over a list of tensor tuples, cut into 8192-element chunks."""

    def __init__(self, columns: list[list[torch.Tensor]], dev: torch.device):
        chunk = _lib.load().cnx_ema_chunk_elems()
        first = columns[0]
        numel = np.array([t.numel() for t in first], dtype=np.int64)
        prefix = np.zeros(len(first) + 1, dtype=np.int64)
        np.cumsum((numel + chunk - 1) // chunk, out=prefix[1:])
        host = np.concatenate([numel, prefix] +
                              [np.array([t.data_ptr() for t in col], dtype=np.int64) for col in columns])
        self._pinned = torch.from_numpy(host).pin_memory()        # kept alive with the table
        self._refs = [list(col) for col in columns]               # the tensors the pointers name
        self.flat = self._pinned.to(dev, non_blocking=True)
        T = len(first)
        self.T, self.total = T, int(prefix[-1])
        self.numels = self.flat[:T]
        self.prefix = self.flat[T:2 * T + 1]
        self.cols = [self.flat[2 * T + 1 + i * T: 2 * T + 1 + (i + 1) * T] for i in range(len(columns))]
        self._staged = []

    def update_col(self, i: int, tensors: list[torch.Tensor]) -> None:
        """Re-point one column (gradients are re-allocated by autograd after zero_grad())."""
        host = torch.from_numpy(np.array([t.data_ptr() for t in tensors], dtype=np.int64)).pin_memory()
        self.cols[i].copy_(host, non_blocking=True)
        self._refs[i] = list(tensors)
        self._staged = self._staged[-3:] + [host]      # keep the source alive until the copy ran


class FusedOptimStep:
    def __init__(self, model: torch.nn.Module, optimizer: torch.optim.Optimizer, scaler=None,
                 ema=None, max_norm: float = 1.0):
        if not isinstance(optimizer, torch.optim.AdamW):
            raise _lib.CnxError("FusedOptimStep implements torch.optim.AdamW (the reference's "
                                "optimizer, train_advanced_v4.py:506-510) only")
        for g in optimizer.param_groups:
            if g.get("amsgrad") or g.get("maximize") or g.get("differentiable"):
                raise _lib.CnxError("FusedOptimStep: amsgrad / maximize / differentiable are not supported")
        self.model, self.optimizer, self.scaler, self.ema = model, optimizer, scaler, ema
        self.max_norm = float(max_norm)
        self._names = {p: n for n, p in model.named_parameters()}
        self._state = None          # device float[8], see include/cnx_b200.h
        self._tables = None
        self._key = None
        self._gkey = None

    # ------------------------------------------------------------------------------ public
    @property
    def total_norm(self) -> torch.Tensor:
        """Unscaled global gradient norm of the last step (device scalar, like clip_grad_norm_)."""
        return self._state[0]

    @property
    def found_inf(self) -> torch.Tensor:
        return self._state[2]

    def step(self) -> None:
        opt = self.optimizer
        groups = []
        for g in opt.param_groups:
            ps = [p for p in g["params"] if p.grad is not None]
            if ps:
                groups.append((g, ps))
        if not groups:
            return
        dev = groups[0][1][0].device
        if dev.type != "cuda":
            raise _lib.CnxError("FusedOptimStep: parameters must live on a CUDA device (no CPU path)")
        if self._state is None:
            self._state = torch.zeros(8, device=dev, dtype=torch.float32)
        self._ensure_state(groups, dev)
        key = self._pointer_key(groups)
        gkey = tuple(p.grad.data_ptr() for _, ps in groups for p in ps)
        if key != self._key:
            self._build(groups, dev)
            # _build re-homes the step counters: take the key of what the tables now point at
            self._key, self._gkey = self._pointer_key(groups), gkey
        all_t, group_ts, rest_t, partials, steps_flat = self._tables
        if gkey != self._gkey:                   # same parameters, gradients moved: pointers only
            all_t.update_col(0, [p.grad for _, ps in groups for p in ps])
            for (_, ps), t in zip(groups, group_ts):
                t.update_col(1, [p.grad for p in ps])
            self._gkey = gkey
        stream = torch.cuda.current_stream(dev).cuda_stream
        scale_ptr = 0
        if self.scaler is not None and self.scaler.is_enabled():
            scale, growth_tracker = self.scaler._check_scale_growth_tracker("step")
            scale_ptr = scale.data_ptr()
        _lib.call("cnx_optim_grad_norm", all_t.cols[0].data_ptr(), all_t.numels.data_ptr(),
                  all_t.prefix.data_ptr(), all_t.T, all_t.total, scale_ptr, self.max_norm,
                  partials.data_ptr(), self._state.data_ptr(), stream)
        omd = d = 0.0
        if self.ema is not None:
            omd = float(np.float32(1.0 - self.ema.decay))      # float64 subtraction first (:79)
            d = float(np.float32(self.ema.decay))
        for (g, _), t in zip(groups, group_ts):
            b1, b2 = g["betas"]
            shadow_ptr = t.cols[4].data_ptr() if self.ema is not None else 0
            _lib.call("cnx_optim_adamw_ema", t.cols[0].data_ptr(), t.cols[1].data_ptr(),
                      t.cols[2].data_ptr(), t.cols[3].data_ptr(), shadow_ptr, t.numels.data_ptr(),
                      t.prefix.data_ptr(), t.T, t.total, float(g["lr"]), float(b1), float(b2),
                      float(g["eps"]), float(g["weight_decay"]), omd, d, self._state.data_ptr(),
                      stream)
        if rest_t is not None:
            # requires_grad parameters that never receive a gradient (the reference's unused aux
            # heads, :146-149): AdamW skips them, EMA.update does not (:76-80)
            _lib.call("cnx_ema_lerp_multi", rest_t.cols[0].data_ptr(), rest_t.cols[1].data_ptr(),
                      rest_t.numels.data_ptr(), rest_t.prefix.data_ptr(), rest_t.T, rest_t.total,
                      omd, d, stream)
        steps_flat.copy_(self._state[3].expand_as(steps_flat))     # torch keeps a count per tensor
        opt._opt_called = True          # what Optimizer.step() records for the LR scheduler's check
        # the parameters were written through raw pointers: torch's version counters did not move,
        # so the blocks' folded-weight caches are invalidated explicitly
        notify_weights_changed()
        if scale_ptr:
            torch._amp_update_scale_(scale, growth_tracker, self._state[2:3],
                                     self.scaler.get_growth_factor(),
                                     self.scaler.get_backoff_factor(),
                                     self.scaler.get_growth_interval())

    # ----------------------------------------------------------------------------- helpers
    def _pointer_key(self, groups):
        """Every pointer the tables hold: parameters, both moments, the step counters and the EMA
        shadows (optimizer.load_state_dict(), ema.register() or a shadow loaded from a checkpoint
        re-allocate them behind a live FusedOptimStep; the tables must follow)."""
        st = self.optimizer.state
        key = tuple((p.data_ptr(), st[p]["exp_avg"].data_ptr(), st[p]["exp_avg_sq"].data_ptr(),
                     st[p]["step"].data_ptr() if torch.is_tensor(st[p]["step"]) else -1)
                    for _, ps in groups for p in ps)
        if self.ema is not None:
            key += tuple(s.data_ptr() for s in self.ema.shadow.values())
        return key

    def _ensure_state(self, groups, dev):
        opt = self.optimizer
        for _, ps in groups:
            for p in ps:
                if p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise _lib.CnxError("FusedOptimStep: parameters and gradients must be float32")
                if not _dense(p):
                    raise _lib.CnxError("FusedOptimStep: parameters must be dense "
                                        f"({self._names.get(p, '?')}: strides {p.stride()})")
                if not _same_layout(p, p.grad):
                    # e.g. a channels_last convolution weight whose gradient came back contiguous:
                    # element i of the gradient must be element i of the parameter in memory
                    p.grad = torch.empty_like(p, memory_format=torch.preserve_format).copy_(p.grad)
                st = opt.state[p]
                if "exp_avg" not in st:
                    st["step"] = torch.zeros((), device=dev, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    self._key = None

    def _build(self, groups, dev):
        opt = self.optimizer
        flat_p = [p for _, ps in groups for p in ps]
        # one step counter per tensor for torch (views of one buffer), one shared count in _state[3]
        steps_flat = torch.zeros(len(flat_p), device=dev, dtype=torch.float32)
        old = opt.state[flat_p[0]]["step"]
        self._state[3] = old.to(dev, torch.float32) if torch.is_tensor(old) else float(old)
        steps_flat.fill_(float(self._state[3].item()))      # table (re)build only: sync is fine here
        for i, p in enumerate(flat_p):
            opt.state[p]["step"] = steps_flat[i]
        all_t = _Table([[p.grad for p in flat_p]], dev)
        group_ts = []
        for _, ps in groups:
            cols = [ps, [p.grad for p in ps], [opt.state[p]["exp_avg"] for p in ps],
                    [opt.state[p]["exp_avg_sq"] for p in ps]]
            if self.ema is not None:
                cols.append([self._shadow_of(p) for p in ps])
            group_ts.append(_Table(cols, dev))
        rest_t = None
        if self.ema is not None:
            have = set(flat_p)
            rest = [p for p in self.model.parameters() if p.requires_grad and p not in have]
            if rest:
                rest_t = _Table([rest, [self._shadow_of(p) for p in rest]], dev)
        partials = torch.empty(all_t.total, device=dev, dtype=torch.float32)
        self._tables = (all_t, group_ts, rest_t, partials, steps_flat)

    def _shadow_of(self, p):
        name = self._names.get(p)
        if name is None or name not in self.ema.shadow:
            raise _lib.CnxError("FusedOptimStep: a parameter with a gradient has no EMA shadow "
                                "(EMA.register() runs over model.named_parameters(), :71-73)")
        s = self.ema.shadow[name]
        if s.dtype != torch.float32 or s.stride() != p.stride():
            raise _lib.CnxError(f"FusedOptimStep: EMA shadow of {name} must be float32 and laid out "
                                "like its parameter")
        return s
