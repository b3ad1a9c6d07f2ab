"""GraphedTrainStep: the body of the reference's hot loop as ONE CUDA-graph launch.

    step = b200ddp.GraphedTrainStep(model, loss_fn, optimizer)
    for inputs, labels in loader:                       # codes/task2/model.py:49-63
        loss = step(inputs, labels)                     # forward, backward, sync, optimizer.step()
        loss_total += loss.item()

The lab's models are small: one step of ResNet-18 on 32x32 images enqueues
~300 kernels that take 2.4 ms on a B200 but 4 ms to launch from Python
(profiles/r01_launch_list_bench_n1.csv), and at 8 GPUs every step waits for the
slowest of eight such hosts.  Replaying the step from a graph removes the host
from the loop.  The kernels of this library take their tensor tables by value
and keep their barrier generations on the device, so they replay correctly.

Semantics are those of the eager loop: the first `eager_steps` calls run
eagerly (they create the momentum buffers with torch's first-step rule and let
cuDNN settle); the next call captures the steady-state step and every call from
then on replays it.  Inputs must keep the shape and dtype of the captured call;
a different shape falls back to the eager path for that call.
"""
from __future__ import annotations

import torch

from .dist_utils import allreduce_average_gradients


class GraphedTrainStep:
    def __init__(self, model, loss_fn, optimizer, sync=allreduce_average_gradients, eager_steps=3):
        self.model, self.loss_fn, self.optimizer, self.sync = model, loss_fn, optimizer, sync
        self.eager_steps = max(1, int(eager_steps))     # at least one: momentum buffers must exist before capture
        self.calls = 0
        self.graph = None
        self.static_in = self.static_target = self.static_loss = None
        self.capture_error = None

    # the reference's loop body, verbatim order (codes/task2/model.py:54-62)
    def _eager(self, inputs, labels):
        outputs = self.model(inputs)
        loss = self.loss_fn(outputs, labels)
        self.optimizer.zero_grad()
        loss.backward()
        self.sync(self.model)
        self.optimizer.step()
        return loss.detach()

    def _capture(self, inputs, labels):
        dev = next(self.model.parameters()).device
        self.static_in = torch.empty(inputs.shape, dtype=inputs.dtype, device=dev)
        self.static_target = torch.empty(labels.shape, dtype=labels.dtype, device=dev)
        self.optimizer.zero_grad(set_to_none=True)      # gradients are re-created inside the graph's memory pool
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            outputs = self.model(self.static_in)
            loss = self.loss_fn(outputs, self.static_target)
            loss.backward()
            self.sync(self.model)
            self.optimizer.step()
            static_loss = loss.detach()
        self.graph, self.static_loss = graph, static_loss
        self._params = [p for p in self.model.parameters() if p.grad is not None]
        self._static_grads = [p.grad for p in self._params]

    def __call__(self, inputs, labels):
        self.calls += 1
        if self.graph is None and self.capture_error is None and self.calls > self.eager_steps:
            try:
                self._capture(inputs, labels)
            except Exception as e:                      # keep training; the caller can inspect capture_error
                self.capture_error = e
                self.graph = None
                torch.cuda.synchronize()
        if self.graph is None or inputs.shape != self.static_in.shape or labels.shape != self.static_target.shape:
            dev = next(self.model.parameters()).device
            return self._eager(inputs.to(dev, non_blocking=True), labels.to(dev, non_blocking=True))
        self.static_in.copy_(inputs, non_blocking=True)
        self.static_target.copy_(labels, non_blocking=True)
        self.graph.replay()
        for p, g in zip(self._params, self._static_grads):   # an eager fallback call may have re-pointed p.grad
            if p.grad is not g:
                p.grad = g
        return self.static_loss
