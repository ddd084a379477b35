"""Optimizers whose update is fused into the gradient reducer.

``FusedSGD`` / ``FusedAdam`` are ``torch.optim.Optimizer`` subclasses: ``param_groups``, LR
schedulers (the reference uses ``StepLR``, pytorch/distributed_data_parallel.py:97,101),
``state_dict()`` / ``load_state_dict()`` in torch's format (``momentum_buffer`` / ``exp_avg`` /
``exp_avg_sq`` per parameter) all work.  The arithmetic is torch's
(``torch/optim/sgd.py``: ``g += wd*p; buf = mu*buf + (1-damp)*g; p -= lr*buf``; Adam with bias
correction), Keras' epsilon placement is available for the TF2 examples
(tensorflow2/mnist_single.py:78).

When the model is wrapped (``wrap`` / ``DataParallel``), the update of a bucket runs inside the
bucket's all-reduce kernel, and ``step()`` only joins the communication stream, re-packs the
bf16 weight operands and advances the step counter.  ``lr`` etc. are read from ``param_groups``
every step (through a small device-side hyper-parameter block), never baked in.

One optimizer object can drive several replicas (the single-process multi-GPU engine): state
arenas, hyper-parameter blocks and weight-pack tables then exist once per GPU.
"""
from __future__ import annotations

import contextlib
import math

import torch

from .. import _native as nv

OPT_SGD, OPT_ADAM, OPT_ADAM_KERAS = 1, 2, 3


def _on(device):
    return torch.cuda.device(device) if device.type == "cuda" else contextlib.nullcontext()


class _FusedBase(torch.optim.Optimizer):
    opt_code = 0
    state_names = ()
    hyper_changes_every_step = False

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        if len(self.param_groups) != 1:
            raise ValueError("the fused optimizers support a single parameter group")
        self._step = 0
        self._reducer = None
        self._arenas = []          # one per replica this optimizer drives
        self._states = []          # per replica: {state name: flat fp32 tensor}
        self._hyper_host = None
        self._hyper_guards = []
        self._hyper_devs = []
        self._pack_caches = []
        self._own_packed = None
        arena = getattr(self._params()[0], "_dtdl_arena", None)
        if arena is not None and getattr(arena, "reducer", None) is not None:
            self.attach(arena.reducer)  # optimizer built after wrap(), as the reference does

    # ------------------------------------------------------------------ wiring
    def _params(self):
        return [p for p in self.param_groups[0]["params"] if p.requires_grad]

    def attach(self, reducer):
        """Called by ``wrap`` / ``DataParallel``: the reducer owns the arenas this optimizer updates."""
        self._reducer = reducer
        arenas = getattr(reducer, "arenas", None) or [reducer.arena]
        self._bind(arenas)
        reducer.optimizer = self

    def _bind(self, arenas):
        self._arenas = list(arenas)
        self._states, self._hyper_devs, self._pack_caches = [], [], []
        self._hyper_host = torch.zeros(16, dtype=torch.float32)
        if self._arenas[0].device.type == "cuda":
            self._hyper_host = self._hyper_host.pin_memory()
        for a in self._arenas:
            multi_proc = (a.peer is not None and a.peer.world > 1
                          and getattr(a.peer, "_local_group", None) is None)
            with _on(a.device):
                self._states.append({n: a.new_state(symmetric=multi_proc) for n in self.state_names})
                self._hyper_devs.append(torch.zeros(16, dtype=torch.float32, device=a.device)
                                        if a.device.type == "cuda" else None)
            self._pack_caches.append({})
        self.refresh_hyper()

    def _ensure_arena(self):
        if not self._arenas:
            from .arena import ParamArena
            params = self._params()
            arena = getattr(params[0], "_dtdl_arena", None)
            if arena is None:
                arena = ParamArena(params, params[0].device)
            self._bind([arena])

    # single-replica aliases used by the multi-process reducer
    @property
    def _arena(self):
        return self._arenas[0] if self._arenas else None

    @property
    def _state_flat(self):
        return self._states[0]

    @property
    def _hyper_dev(self):
        return self._hyper_devs[0]

    def _state_on(self, g):
        st = self._states[g]
        return st[self.state_names[0]], (st[self.state_names[1]] if len(self.state_names) > 1 else None)

    def _hyper_on(self, g):
        return self._hyper_devs[g]

    # ------------------------------------------------------------------ hyper-parameters
    def _hyper_values(self, world):
        raise NotImplementedError

    def _world(self):
        r = self._reducer
        return getattr(r, "world", 1) if r is not None else 1

    def refresh_hyper(self, enqueue_copy=True):
        """Writes the current ``param_groups`` values into the pinned block and (optionally)
        enqueues its copy to the device(s).  A captured CUDA graph re-reads the pinned block on
        every replay, so calling this with ``enqueue_copy=False`` before a replay is enough."""
        vals = self._hyper_values(self._world())
        new = torch.tensor(vals, dtype=torch.float32)
        # Under stream capture nothing executes and the CUDA runtime refuses event waits from the host
        # (they would invalidate the capture): the capturing caller (engine.TrainStep._capture,
        # DataParallelStep) refreshes the block and drains earlier work BEFORE it starts the capture.
        capturing = self._hyper_host.is_pinned() and torch.cuda.is_current_stream_capturing()
        if not torch.equal(self._hyper_host[: len(vals)], new):
            # The block is about to change while an earlier step's copy of it (enqueued below, or a
            # node of a replayed graph) may not have run yet: that step must not see the new values.
            if not capturing:
                for ev in self._hyper_guards:
                    ev.synchronize()
            self._hyper_host[: len(vals)] = new
        if not capturing:
            self._hyper_guards = []
        if enqueue_copy:
            for h in self._hyper_devs:
                if h is not None:
                    with torch.cuda.device(h.device):
                        h.copy_(self._hyper_host, non_blocking=True)
                        if not capturing:       # an event recorded into a capture cannot be waited for
                            ev = torch.cuda.Event()
                            ev.record()
                            self._hyper_guards.append(ev)

    def guard_hyper(self, event):
        """``event`` completes after a consumer of the pinned block (a graph replay) has read it."""
        self._hyper_guards.append(event)

    # ------------------------------------------------------------------ step
    def zero_grad(self, set_to_none: bool = True):
        """No memset for kernel-written gradients: the next backward overwrites them (see
        ops.functional.grad_buffer).  Gradients produced by stock torch ops are dropped."""
        plist = self._params() if len(self._arenas) <= 1 else [p for a in self._arenas for p in a.params]
        for p in plist:
            if not getattr(p, "_dtdl_native", False):
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
            self._local_update_on(0)
        self._step += 1
        self._repack()
        return loss

    def _local_update_on(self, g):
        a = self._arenas[g]
        if a.device.type == "cuda":
            m, v = self._state_on(g)
            with torch.cuda.device(a.device):
                nv.call("local_update", nv.ptr(a.param_flat), nv.ptr(a.grad_flat), nv.ptr(m), nv.ptr(v),
                        nv.ptr(self._hyper_devs[g]), a.numel, self.opt_code,
                        torch.cuda.current_stream(a.device).cuda_stream)
        else:
            self._update_torch(a.grad_flat, 0, a.numel)

    def _local_update(self):
        self._local_update_on(0)

    def _update_torch(self, grad_flat, off, n):
        raise NotImplementedError

    def _repack(self):
        """bf16 operand packs of conv / linear weights follow the fp32 master weights."""
        if self._arenas[0].device.type != "cuda":
            return
        from ..ops.modules import pack_weights
        r = self._reducer
        if r is not None and hasattr(r, "repack_all"):
            r.repack_all(self._pack_caches)
            return
        if r is not None and getattr(r, "packs_current", False):
            return    # every bucket re-packed its own layers on the communication stream
        mods = r.packed_modules if r is not None else self._own_packed
        if mods:
            pack_weights(mods, self._pack_caches[0])

    def register_packed_modules(self, module):
        """Bare (unwrapped) model: tell the optimizer which layers need re-packing after a step."""
        from ..ops.modules import packed_layers
        self._own_packed = packed_layers(module)

    # ------------------------------------------------------------------ torch-format state
    def _gather_state(self):
        """Two-shot fused updates shard the optimizer state by rank; make replica 0's copy whole."""
        red = self._reducer
        if red is None or self._step == 0 or not getattr(red, "state_sharded", False):
            return
        a0 = self._arenas[0]
        if len(self._arenas) > 1:        # single process: copy each owner's slice over
            W = len(self._arenas)
            for name in self.state_names:
                for b in a0.buckets:
                    if not a0.peer.choose_two_shot(b.numel * 4):
                        continue
                    n4 = b.numel // 4
                    per = (n4 + W - 1) // W
                    for r in range(1, W):
                        lo, hi = b.offset + 4 * per * r, b.offset + 4 * min(per * (r + 1), n4)
                        if hi > lo:
                            self._states[0][name][lo:hi].copy_(self._states[r][name][lo:hi])
        elif getattr(red, "peer", None) is not None and red.world > 1:
            for t in self._states[0].values():
                buf = getattr(t, "_dtdl_symm", None)
                if buf is None:
                    continue
                for b in a0.buckets:
                    if red.peer.choose_two_shot(b.numel * 4) or getattr(red, "grad_bf16", False):
                        red.peer.allgather(buf, b.offset * 4, b.numel * 4)
        if a0.device.type == "cuda":
            torch.cuda.synchronize(a0.device)

    def state_dict(self):
        self._ensure_arena()
        self._gather_state()
        params = self.param_groups[0]["params"]
        state = {}
        if self._step > 0:
            for i, p in enumerate(params):
                if not p.requires_grad:
                    continue
                entry = {k: self._arena.flat_view_like(t, p).clone() for k, t in self._states[0].items()}
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
        for st in self._states:
            for t in st.values():
                t.zero_()
        for i, entry in sd.get("state", {}).items():
            p = params[int(i)]
            off = self._arenas[0].offsets[id(p)]
            for st in self._states:
                for k, t in st.items():
                    if k in entry:
                        t[off: off + p.numel()].copy_(entry[k].reshape(-1))
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
        self.keras = keras
        if keras:
            self.opt_code = OPT_ADAM_KERAS
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

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
