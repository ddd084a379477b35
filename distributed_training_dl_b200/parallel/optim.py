"""Optimizers whose update is fused into the gradient reducer.

``FusedSGD`` / ``FusedAdam`` are ``torch.optim.Optimizer`` subclasses: ``param_groups``, LR
schedulers (the reference uses ``StepLR``, pytorch/distributed_data_parallel.py:97,101),
``state_dict()`` / ``load_state_dict()`` in torch's format (``momentum_buffer`` / ``exp_avg`` /
``exp_avg_sq`` per parameter) all work.  The arithmetic is torch's
(``torch/optim/sgd.py``: ``g += wd*p; buf = mu*buf + (1-damp)*g; p -= lr*buf``; Adam with bias
correction), Keras' epsilon placement is available for the TF2 examples
(tensorflow2/mnist_single.py:78).

When the model is wrapped (:func:`..wrap.wrap`), the update of a bucket runs inside the
bucket's all-reduce kernel during backward, and ``step()`` only joins the communication stream,
re-packs the bf16 weight operands and advances the step counter.  ``lr`` etc. are read from
``param_groups`` every step (through a small device-side hyper-parameter block), never baked in.
"""
from __future__ import annotations

import math

import torch

from .. import _native as nv
from ..ops import functional as fn

OPT_SGD, OPT_ADAM, OPT_ADAM_KERAS = 1, 2, 3


class _FusedBase(torch.optim.Optimizer):
    opt_code = 0
    state_names = ()

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        if len(self.param_groups) != 1:
            raise ValueError("the fused optimizers support a single parameter group")
        self._step = 0
        self._reducer = None
        self._arena = None
        self._state_flat = {}
        self._hyper_host = None
        self._hyper_dev = None
        self._pack_cache = {}
        arena = getattr(self._params()[0], "_dtdl_arena", None)
        if arena is not None and getattr(arena, "reducer", None) is not None:
            self.attach(arena.reducer)  # optimizer built after wrap(), as the reference does

    # ------------------------------------------------------------------ wiring
    def _params(self):
        return [p for p in self.param_groups[0]["params"] if p.requires_grad]

    def attach(self, reducer):
        """Called by ``wrap``: the reducer owns the arenas this optimizer updates."""
        self._reducer = reducer
        self._bind_arena(reducer.arena)
        reducer.optimizer = self

    def _bind_arena(self, arena):
        self._arena = arena
        dev = arena.device
        symmetric = arena.peer is not None and arena.peer.world > 1
        for name in self.state_names:
            self._state_flat[name] = arena.new_state(symmetric=symmetric)
        self._hyper_host = torch.zeros(16, dtype=torch.float32)
        if dev.type == "cuda":
            self._hyper_host = self._hyper_host.pin_memory()
            self._hyper_dev = torch.zeros(16, dtype=torch.float32, device=dev)
        self.refresh_hyper()

    def _ensure_arena(self):
        if self._arena is None:
            from .arena import ParamArena
            params = self._params()
            arena = getattr(params[0], "_dtdl_arena", None)
            if arena is None:
                arena = ParamArena(params, params[0].device)
            self._bind_arena(arena)

    # ------------------------------------------------------------------ hyper-parameters
    def _hyper_values(self, world):
        raise NotImplementedError

    def refresh_hyper(self, enqueue_copy=True):
        """Writes the current ``param_groups`` values into the pinned block and (optionally)
        enqueues its copy to the device.  A captured CUDA graph re-reads the pinned block on
        every replay, so calling this with ``enqueue_copy=False`` before a replay is enough."""
        world = self._reducer.world if self._reducer is not None else 1
        vals = self._hyper_values(world)
        self._hyper_host[: len(vals)] = torch.tensor(vals, dtype=torch.float32)
        if self._hyper_dev is not None and enqueue_copy:
            self._hyper_dev.copy_(self._hyper_host, non_blocking=True)

    # ------------------------------------------------------------------ step
    def zero_grad(self, set_to_none: bool = True):
        """No memset for kernel-written gradients: the next backward overwrites them (see
        ops.functional.grad_buffer).  Gradients produced by stock torch ops are dropped."""
        for p in self._params():
            if getattr(p, "_dtdl_native", False):
                p._dtdl_fresh = True
            else:
                p.grad = None
                p._dtdl_fresh = True

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self._ensure_arena()
        if self._reducer is not None:
            self._reducer.finish_step(fused=True)
        else:
            for p in self._params():  # bare model: gather torch-produced gradients
                g = p._dtdl_grad
                if p.grad is not None and p.grad.data_ptr() != g.data_ptr():
                    g.copy_(p.grad)
                    p.grad = g
            self.refresh_hyper()
            self._local_update()
        self._step += 1
        self._repack()
        return loss

    def _local_update(self):
        a = self._arena
        if a.device.type == "cuda":
            m = self._state_flat[self.state_names[0]]
            v = self._state_flat[self.state_names[1]] if len(self.state_names) > 1 else None
            nv.call("local_update", nv.ptr(a.param_flat), nv.ptr(a.grad_flat), nv.ptr(m), nv.ptr(v),
                    nv.ptr(self._hyper_dev), a.numel, self.opt_code, nv.stream_ptr())
        else:
            self._update_torch(a.grad_flat, 0, a.numel)

    def _update_torch(self, grad_flat, off, n):
        raise NotImplementedError

    def _repack(self):
        """bf16 operand packs of conv / linear weights follow the fp32 master weights."""
        if self._arena.device.type != "cuda":
            return
        mods = self._reducer.packed_modules if self._reducer is not None else self._pack_cache.get("mods")
        if mods:
            from ..ops.modules import pack_weights
            pack_weights(mods, self._pack_cache)

    def register_packed_modules(self, module):
        from ..ops.modules import packed_layers
        self._pack_cache = {"mods": packed_layers(module)}

    # ------------------------------------------------------------------ torch-format state
    def _gather_state(self):
        red = self._reducer
        if red is not None and red.world > 1 and red.peer is not None and self._step > 0:
            for t in self._state_flat.values():
                buf = getattr(t, "_dtdl_symm", None)
                if buf is not None and red.state_sharded:
                    red.peer.allgather(buf, 0, self._arena.numel * 4)
            torch.cuda.synchronize(self._arena.device)

    def state_dict(self):
        self._ensure_arena()
        self._gather_state()
        params = self.param_groups[0]["params"]
        state = {}
        if self._step > 0:
            for i, p in enumerate(params):
                if not p.requires_grad:
                    continue
                entry = {k: self._arena.flat_view_like(t, p).clone() for k, t in self._state_flat.items()}
                entry.update(self._extra_state())
                state[i] = entry
        group = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        group["params"] = list(range(len(params)))
        return {"state": state, "param_groups": [group], "dtdl_step": self._step}

    def load_state_dict(self, sd):
        self._ensure_arena()
        params = self.param_groups[0]["params"]
        g = dict(sd["param_groups"][0])
        g.pop("params", None)
        self.param_groups[0].update(g)
        for t in self._state_flat.values():
            t.zero_()
        for i, entry in sd.get("state", {}).items():
            p = params[int(i)]
            for k, t in self._state_flat.items():
                if k in entry:
                    self._arena.flat_view_like(t, p).copy_(entry[k])
        steps = [int(e["step"]) for e in sd.get("state", {}).values() if "step" in e]
        self._step = int(sd.get("dtdl_step", max(steps) if steps else (1 if sd.get("state") else 0)))
        self.refresh_hyper()

    def _extra_state(self):
        return {}


class FusedSGD(_FusedBase):
    """SGD with momentum and weight decay (pytorch/single_gpu.py:71-72)."""
    opt_code = OPT_SGD
    state_names = ("momentum_buffer",)

    def __init__(self, params, lr=0.1, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if nesterov:
            raise ValueError("nesterov momentum is not implemented (the reference does not use it)")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening,
                                      weight_decay=weight_decay, nesterov=False))

    def _hyper_values(self, world):
        g = self.param_groups[0]
        # torch's first step sets buf = g (no dampening); with buf initialised to zero that is
        # mu*0 + (1-damp)*g, identical when damp == 0, and handled by zeroing damp on step 0.
        damp = g["dampening"] if self._step > 0 else 0.0
        return [g["lr"], g["momentum"], g["weight_decay"], 1.0 / world, 0, 0, 0, 1, 1, damp]

    def _update_torch(self, grad_flat, off, n):
        g = self.param_groups[0]
        a = self._arena
        p = a.param_flat[off: off + n]
        m = self._state_flat["momentum_buffer"][off: off + n]
        d = grad_flat[off: off + n]
        if g["weight_decay"]:
            d = d.add(p, alpha=g["weight_decay"])
        damp = g["dampening"] if self._step > 0 else 0.0
        m.mul_(g["momentum"]).add_(d, alpha=1.0 - damp)
        p.add_(m, alpha=-g["lr"])


class FusedAdam(_FusedBase):
    """Adam.  ``keras=True`` uses Keras' epsilon placement and default eps 1e-7
    (tensorflow2/mnist_single.py:78); otherwise torch / Chainer's (chainer/train_mnist.py:69)."""
    opt_code = OPT_ADAM
    state_names = ("exp_avg", "exp_avg_sq")
    hyper_changes_every_step = True  # bias corrections

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=None, weight_decay=0.0, keras=False):
        eps = (1e-7 if keras else 1e-8) if eps is None else eps
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.keras = keras
        if keras:
            self.opt_code = OPT_ADAM_KERAS

    def _hyper_values(self, world):
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        t = self._step + 1
        return [g["lr"], 0.0, g["weight_decay"], 1.0 / world, b1, b2, g["eps"], 1 - b1 ** t,
                math.sqrt(1 - b2 ** t), 0.0]

    def _extra_state(self):
        return {"step": torch.tensor(float(self._step))}

    def _update_torch(self, grad_flat, off, n):
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        t = self._step + 1
        a = self._arena
        p = a.param_flat[off: off + n]
        m = self._state_flat["exp_avg"][off: off + n]
        v = self._state_flat["exp_avg_sq"][off: off + n]
        d = grad_flat[off: off + n]
        if g["weight_decay"]:
            d = d.add(p, alpha=g["weight_decay"])
        m.mul_(b1).add_(d, alpha=1 - b1)
        v.mul_(b2).addcmul_(d, d, value=1 - b2)
        bc1, bc2s = 1 - b1 ** t, math.sqrt(1 - b2 ** t)
        if self.keras:
            p.addcdiv_(m, v.sqrt().add_(g["eps"]), value=-g["lr"] * bc2s / bc1)
        else:
            p.addcdiv_(m, (v.sqrt() / bc2s).add_(g["eps"]), value=-g["lr"] / bc1)
