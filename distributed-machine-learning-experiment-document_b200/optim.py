"""SGD: torch.optim.SGD's constructor, state layout and update rule, applied by
one multi-tensor kernel (K3).

Stands in for `torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)` +
`optimizer.step()` (codes/task2/model.py:131,62; codes/task3/model.py:118);
arithmetic of torch/optim/sgd.py:322-380 including the first-step rule
(`momentum_buffer = clone(grad)`, no dampening).  Momentum lives in
`optimizer.state[p]["momentum_buffer"]` as in torch (sgd.py:144-150), so
state_dict() / load_state_dict() keep working.
"""
from __future__ import annotations

import torch
from torch.optim.optimizer import Optimizer

from . import _lib
from .runtime import default_context, opt_args, sgd_args


class SGD(Optimizer):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, *,
                 maximize=False):
        # same argument checks and messages as torch/optim/sgd.py:45-75
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if momentum < 0.0:
            raise ValueError(f"Invalid momentum value: {momentum}")
        if weight_decay < 0.0:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                        nesterov=nesterov, maximize=maximize)
        super().__init__(params, defaults)
        self._plans = {}
        self._engine = None          # set by b200ddp.enable_fused_step
        self._ctx = None             # context of the direct path (None: the process-wide default)
        self._write_grads = True
        self.launches = 0

    @staticmethod
    def _hyper(group):
        return dict(lr=group["lr"], momentum=group["momentum"], dampening=group["dampening"],
                    weight_decay=group["weight_decay"], nesterov=group["nesterov"], maximize=group["maximize"])

    def _momentum_buffers(self, params, momentum):
        """(buffers, fresh): fresh[i] is True where the buffer was just created."""
        if momentum == 0:
            return None, [False] * len(params)
        bufs, fresh = [], []
        for p in params:
            st = self.state[p]
            buf = st.get("momentum_buffer")
            fresh.append(buf is None)
            if buf is None:
                buf = st["momentum_buffer"] = torch.empty_like(p, memory_format=torch.contiguous_format)
            bufs.append(buf)
        return bufs, fresh

    def _direct(self, ctx, params, bufs, hyper, first_step):
        key = (tuple(id(p) for p in params), first_step)
        plan = self._plans.get(key)
        if plan is None:
            plan = self._plans[key] = ctx.plan([p.numel() for p in params])
        for p in params:
            if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous() or not p.grad.is_contiguous():
                raise _lib.B200DDPError("b200ddp.SGD handles contiguous fp32 CUDA parameters and gradients")
            if p.grad.is_sparse:
                raise _lib.B200DDPError("b200ddp.SGD does not support sparse gradients")
        ctx.sgd_step(plan, [p.grad.data_ptr() for p in params], [p.data_ptr() for p in params],
                     [b.data_ptr() for b in bufs] if bufs is not None else None,
                     sgd_args(first_step=first_step, **hyper))
        self.launches += 1

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        eng = self._engine
        if eng is not None and eng.pending and eng.world == 1:
            eng.pending = False      # one rank: the mean is the local gradient, update from p.grad directly
        if eng is not None and eng.pending:
            # fused path: gradients are still local; pack -> reduce -> update in one go
            group = self.param_groups[0]
            hyper = self._hyper(group)
            bufs, fresh = self._momentum_buffers(eng.params, hyper["momentum"])
            if any(fresh) and not all(fresh):
                raise _lib.B200DDPError("fused step needs all momentum buffers created together")
            before = eng.launches
            eng.sync_and_step(bufs, hyper, first_step=bool(fresh and fresh[0]), write_grads=self._write_grads)
            self.launches += eng.launches - before
            eng.pending = False
            return loss
        ctx = self._ctx if self._ctx is not None else default_context()
        for group in self.param_groups:
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            hyper = self._hyper(group)
            bufs, fresh = self._momentum_buffers(params, hyper["momentum"])
            for want in (True, False):           # first-step and steady-state tensors go in separate launches
                sel = [i for i, f in enumerate(fresh) if f == want]
                if sel:
                    self._direct(ctx, [params[i] for i in sel], [bufs[i] for i in sel] if bufs is not None else None,
                                 hyper, first_step=want and hyper["momentum"] != 0)
        return loss


# ---------------------------------------------------------------------------------------------
# The lab's own optimizers (task 1): same constructors, attributes and step()/zero_grad() as
# codes/task1/pytorch/MyOptimizer.py:3-43, one multi-tensor kernel per step instead of 1 to 9
# tensor operations per parameter.
# ---------------------------------------------------------------------------------------------
class BaseOptimizer:
    """codes/task1/pytorch/MyOptimizer.py:3-16."""
    _rule = None

    def __init__(self, params, lr=0.001):
        self.params = list(params)
        self.lr = lr
        self._plans = {}
        self._engine = None          # set by b200ddp.enable_fused_step
        self._write_grads = True
        self.launches = 0

    def step(self):
        raise NotImplementedError

    def zero_grad(self):
        for p in self.params:
            if p.grad is not None:
                p.grad.detach_()
                p.grad.zero_()

    def _hyper(self):
        return dict(lr=self.lr)

    def _states(self, idx):
        return None, None

    def _step(self):
        eng = self._engine
        if eng is not None and eng.pending and eng.world == 1:
            eng.pending = False
        if eng is not None and eng.pending:
            m, v = self._states(range(len(self.params)))
            before = eng.launches
            eng.sync_and_step(m, self._hyper(), first_step=False, write_grads=self._write_grads, rule=self._rule, vels=v)
            self.launches += eng.launches - before
            eng.pending = False
            return
        idx = [i for i, p in enumerate(self.params) if p.grad is not None]     # MyOptimizer.py:22,37
        if not idx:
            return
        ps = [self.params[i] for i in idx]
        for p in ps:
            if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous() or not p.grad.is_contiguous():
                raise _lib.B200DDPError("b200ddp optimizers handle contiguous fp32 CUDA parameters and gradients")
        ctx = default_context()
        key = tuple(idx)
        plan = self._plans.get(key)
        if plan is None:
            plan = self._plans[key] = ctx.plan([p.numel() for p in ps])
        m, v = self._states(idx)
        ctx.opt_step(plan, [p.grad.data_ptr() for p in ps], [p.data_ptr() for p in ps],
                     [t.data_ptr() for t in m] if m is not None else None,
                     [t.data_ptr() for t in v] if v is not None else None,
                     opt_args(self._rule, **self._hyper()))
        self.launches += 1


class GdOptimizer(BaseOptimizer):
    """codes/task1/pytorch/MyOptimizer.py:18-23: p = p - lr * grad."""
    _rule = "gd"

    @torch.no_grad()
    def step(self):
        self._step()


class AdamOptimizer(BaseOptimizer):
    """codes/task1/pytorch/MyOptimizer.py:25-43: Adam without bias correction, state in
    `momentums` / `velocities` (zeros at construction), every tensor operation of the reference's
    expression rounded on its own."""
    _rule = "adam"

    def __init__(self, params, lr=0.001, b1=0.9, b2=0.999, epsilon=1e-8):
        super().__init__(params, lr)
        self.beta1 = b1
        self.beta2 = b2
        self.epsilon = epsilon
        self.momentums = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]
        self.velocities = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]

    def _hyper(self):
        return dict(lr=self.lr, beta1=self.beta1, beta2=self.beta2, eps=self.epsilon)

    def _states(self, idx):
        return [self.momentums[i] for i in idx], [self.velocities[i] for i in idx]

    @torch.no_grad()
    def step(self):
        self._step()
