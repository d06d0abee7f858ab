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
