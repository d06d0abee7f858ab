"""Forward + loss + scaled backward of one fixed-shape micro-batch as ONE CUDA-graph launch.

The EfficientNet-B2 step of the older reference loop
(/root/reference/previous/V3-efb2/train_advanced_v3.0.py:96-113: autocast forward, criterion,
`scaler.scale(loss).backward()`, batch 48 at 260x260) is ~1 500 small kernels; enqueued one by one
from Python the host, not the B200, sets the step time.  Every kernel of libcnx_b200 is launched on
the caller's stream with no host synchronisation and no library-owned memory, so the whole
micro-step can be recorded once and replayed: `GraphedForwardBackward` follows PyTorch's
whole-network capture recipe (side-stream warm-up, static input / output tensors, gradients living
in the graph's private pool).  The optimizer half of the step stays outside the graph
(`optim_step.FusedOptimStep`: two launches, learning rate free to change every step).  With
`grad_reducer=` the data-parallel gradient all-reduce (ddp.BucketedGradReducer, NCCL) is captured
inside the same graph.

Contract: fixed batch shape; parameters are not re-allocated after capture; after `__call__` the
`.grad` of every parameter that received a gradient is the graph's static buffer holding THIS
micro-batch's gradient (overwritten, not accumulated, by the next call) — so the reference's
accumulate-1 loops map onto it directly, `optimizer.zero_grad()` included.
"""
from __future__ import annotations

import torch


class GraphedForwardBackward:
    def __init__(self, model: torch.nn.Module, criterion, scaler, example_images: torch.Tensor,
                 example_labels: torch.Tensor, amp_dtype: torch.dtype = torch.bfloat16,
                 warmup: int = 3, grad_reducer=None):
        if not example_images.is_cuda:
            raise RuntimeError("GraphedForwardBackward needs CUDA tensors (no CPU path)")
        self.model, self.criterion, self.scaler, self.amp_dtype = model, criterion, scaler, amp_dtype
        # ddp.BucketedGradReducer or None: its bucket all-reduces are issued from the gradient
        # hooks DURING the captured backward and joined at its end, so a replay overlaps NCCL
        # with the remaining backward kernels exactly like the eager loop does
        self.reducer = grad_reducer
        self.images = example_images.clone(memory_format=torch.preserve_format)
        self.labels = example_labels.clone()
        self.params = [p for p in model.parameters() if p.requires_grad]
        side = torch.cuda.Stream(device=example_images.device)
        side.wait_stream(torch.cuda.current_stream(example_images.device))
        with torch.cuda.stream(side):
            for _ in range(warmup):             # lazy initialisation (cuDNN plans, TMA tables, ...)
                for p in self.params:
                    p.grad = None
                self._fwd_bwd()
        torch.cuda.current_stream(example_images.device).wait_stream(side)
        for p in self.params:
            p.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss, self.outputs = self._fwd_bwd()
        self.grads = [p.grad for p in self.params]          # static buffers in the graph's pool

    def _fwd_bwd(self):
        if self.reducer is not None:
            self.reducer.set_sync(True)
        with torch.autocast("cuda", dtype=self.amp_dtype):
            outputs = self.model(self.images)
            if isinstance(outputs, (list, tuple)):
                outputs = outputs[0]
            loss = self.criterion(outputs, self.labels)
        (self.scaler.scale(loss) if self.scaler is not None else loss).backward()
        if self.reducer is not None:
            self.reducer.finish()
        return loss.detach(), outputs.detach()

    def __call__(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Copy the batch into the static buffers, replay, re-attach the static gradients.
        Returns the (static) loss tensor; `self.outputs` holds the logits."""
        if images.shape != self.images.shape or labels.shape != self.labels.shape:
            raise RuntimeError(f"GraphedForwardBackward was captured for {tuple(self.images.shape)} / "
                               f"{tuple(self.labels.shape)}, got {tuple(images.shape)} / "
                               f"{tuple(labels.shape)}")
        if images.data_ptr() != self.images.data_ptr():
            self.images.copy_(images, non_blocking=True)
        if labels.data_ptr() != self.labels.data_ptr():
            self.labels.copy_(labels, non_blocking=True)
        self.graph.replay()
        for p, g in zip(self.params, self.grads):
            p.grad = g
        return self.loss


class GraphedTrainStep:
    """`train_step.TrainStep` with forward + loss + backward of every micro-batch position of the
    accumulation window replayed as ONE CUDA graph each (the reference loop,
    /root/reference/train_advanced_v4.py:215-237, accumulates 2 micro-batches: two graphs).

    The ConvNeXt micro-step is ~500 library launches enqueued from Python in ~12 ms; at BASELINE
    config 2's micro-batch of 128 the GPU needs 22 ms for them and the host stays ahead, at config
    3's 8-GPU share (micro-batch 64) it does not.  Everything between "the mixed batch is in HBM"
    and "the gradients are in .grad" is stream-ordered library / torch work with no host sync, so
    it is recorded once per window position and replayed:

      graph 0      zero the GradWindow accumulators, fold the block weights (from the CURRENT
                   parameters: the fold kernels are inside the graph), forward, soft-target CE,
                   scaled backward; block gradients stay in the window's accumulators
      graph k      same without the fold (reads graph 0's folded tensors), accumulating
      graph N-1    ... and the fold adjoint emits every parameter gradient; with a
                   ddp.BucketedGradReducer the bucket all-reduces issued from its gradient hooks
                   (NCCL) and their join are part of this graph
    The optimizer section (fused unscale + clip + AdamW + EMA, scheduler) stays outside: two
    launches, learning rate free to change.  The graphs share one memory pool (replayed in capture
    order, so activations of graph k are recycled by graph k+1).

    Contract: fixed batch shape; `micro_step(images, targets, batch_idx)` copies its arguments
    into the static input buffers unless they already ARE those buffers (`self.images`,
    `self.targets` — e.g. hand them to `mix.BatchMixUpCutMix(..., out=, targets_out=)`); windows
    always start at position 0 (an odd trailing micro-batch is dropped at the next window start,
    unlike the eager `TrainStep`, which carries it over like the reference does)."""

    def __init__(self, step, images_shape, targets_shape, device, warmup_windows: int = 2):
        from . import block as block_mod
        self.step = step
        self.N = step.cfg.gradient_accumulation_steps
        self.device = torch.device(device)
        self.images = torch.zeros(images_shape, device=self.device, dtype=torch.float32)
        self.targets = torch.full(targets_shape, 1.0 / targets_shape[-1], device=self.device,
                                  dtype=torch.float32)
        self.params = step.params
        if step.window is not None:
            step.window.end()
        cur = torch.cuda.current_stream(self.device)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            # eager windows: lazy initialisation, and the reducer's unused-parameter discovery
            for _ in range(warmup_windows):
                for p in self.params:
                    p.grad = None
                for k in range(self.N):
                    step.forward_backward(self.images, self.targets, k)
                if step.grad_reducer is not None:
                    step.grad_reducer.finish()
                if step.window is not None:
                    step.window.end()
        cur.wait_stream(side)
        torch.cuda.synchronize(self.device)
        for p in self.params:
            p.grad = None
        from . import _lib
        pool = torch.cuda.graph_pool_handle()
        self.graphs, self.losses, self.launches = [], [], []
        try:
            for k in range(self.N):
                block_mod.capture_reuses_fold = k > 0     # graph 0 folds (prefold_blocks), the others reuse
                g = torch.cuda.CUDAGraph()
                n0 = _lib.launch_count
                with torch.cuda.graph(g, pool=pool):
                    loss = step.forward_backward(self.images, self.targets, k)
                    if k == self.N - 1 and step.grad_reducer is not None:
                        step.grad_reducer.finish()
                self.graphs.append(g)
                self.losses.append(loss)
                self.launches.append(_lib.launch_count - n0)    # library kernels inside graph k
        finally:
            block_mod.capture_reuses_fold = False
        if step.window is not None:
            step.window.end()
        self.grads = [p.grad for p in self.params]          # static buffers in the graphs' pool
        for p in self.params:
            p.grad = None

    def micro_step(self, images: torch.Tensor, targets: torch.Tensor, batch_idx: int) -> torch.Tensor:
        k = batch_idx % self.N
        if images.shape != self.images.shape or targets.shape != self.targets.shape:
            raise RuntimeError(f"GraphedTrainStep was captured for {tuple(self.images.shape)} / "
                               f"{tuple(self.targets.shape)}, got {tuple(images.shape)} / "
                               f"{tuple(targets.shape)}")
        if images.data_ptr() != self.images.data_ptr():
            self.images.copy_(images, non_blocking=True)
        if targets.data_ptr() != self.targets.data_ptr():
            self.targets.copy_(targets, non_blocking=True)
        self.graphs[k].replay()
        from . import _lib
        _lib.launch_count += self.launches[k]        # the replay launched the recorded kernels
        if k == self.N - 1:
            for p, g in zip(self.params, self.grads):
                p.grad = g
            if self.step.grad_reducer is not None:
                self.step.grad_reducer.reduced_steps += 1      # joined inside the graph
            self.step.optimizer_section(reduce_grads=False)
        return self.losses[k]

    def close(self) -> None:
        self.step.close()
