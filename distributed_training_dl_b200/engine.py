"""Training-step engine: one call = H2D copy, forward, loss, backward, reduce+update.

``TrainStep(model, optimizer, criterion)`` is the user-facing fast path.  The reference's hot
loop (pytorch/single_gpu.py:93-116) launches ~1500 tiny kernels per step from Python and blocks
on three ``.item()`` calls; here the whole step -- including the side-stream bucket
reductions and the fused optimizer -- is captured ONCE into a CUDA graph and replayed, inputs
are copied from pinned host memory into static device buffers, and loss / accuracy counters
stay on the device until the caller asks.  No tracing compiler is involved: capture records the
hand-written kernels exactly as eager mode launches them.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _native as nv
from .ops import functional as fn


class CrossEntropyLoss(nn.Module):
    """Softmax cross entropy (mean) with the accuracy count folded into the same kernel
    (K-O13 + K-O17; the reference: ``nn.CrossEntropyLoss`` + ``outputs.max(1)`` / ``eq`` /
    ``sum`` -- pytorch/single_gpu.py:70,99,105-108).  ``last_stats`` = [loss, #correct]."""

    def __init__(self):
        super().__init__()
        self.last_stats = None

    def forward(self, logits, targets):
        if not logits.is_cuda:
            loss = F.cross_entropy(logits, targets)
            self.last_stats = torch.stack([loss.detach(), (logits.detach().argmax(1) == targets).sum().float()])
            return loss
        loss, stats = fn.softmax_cross_entropy(logits, targets)
        self.last_stats = stats
        return loss


class TrainStep:
    """Callable training step.

    >>> step = TrainStep(net, opt, criterion)           # net = dtdl.wrap(model)
    >>> loss = step(x_pinned_cpu, y_pinned_cpu)         # device scalar; .item() when needed

    ``graph=True`` captures after ``warmup`` eager iterations (which also train).  Batch shape
    must stay fixed once captured; a differently shaped batch falls back to eager for that call.
    """

    def __init__(self, model, optimizer, criterion=None, graph=True, warmup=3):
        self.model, self.optimizer = model, optimizer
        self.criterion = criterion if criterion is not None else CrossEntropyLoss()
        self.device = next(model.parameters()).device
        self.use_graph = graph and self.device.type == "cuda"
        self.warmup = warmup
        self._slots = {}
        self._packed = None
        self._done = None
        self._stats = None
        self.launches_per_step = None

    # ------------------------------------------------------------------
    def _eager(self, x, t):
        out = self.model(x)
        loss = self.criterion(out, t)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        # detached: a caller that keeps the returned loss must not keep this step's autograd graph alive
        # (its gradient-accumulation nodes are bound to the stream of THIS step; a later capture that
        # reused them would depend on uncaptured work)
        return loss.detach(), getattr(self.criterion, "last_stats", None)

    def _slot(self, x, t):
        """Static buffers + graph per batch shape (a smaller last batch gets its own slot and
        simply stays eager until it has been seen ``warmup`` times)."""
        key = (tuple(x.shape), tuple(t.shape))
        slot = self._slots.get(key)
        if slot is None:
            slot = {"sx": torch.empty(x.shape, dtype=x.dtype, device=self.device),
                    "st": torch.empty(t.shape, dtype=t.dtype, device=self.device),
                    "graph": None, "calls": 0, "loss": None, "stats": None}
            self._slots[key] = slot
        slot["sx"].copy_(x, non_blocking=True)
        slot["st"].copy_(t, non_blocking=True)
        return slot

    def _capture(self, slot):
        opt = self.optimizer
        nv.reset_launch_count()
        g = torch.cuda.CUDAGraph()
        step_before = getattr(opt, "_step", None)
        if hasattr(opt, "refresh_hyper"):
            # values that change every step (Adam's bias corrections) go into the pinned block now: the
            # refresh inside the captured iteration then finds nothing to change and nothing to wait for
            opt.refresh_hyper(enqueue_copy=False)
        try:
            with torch.cuda.graph(g):
                loss, stats = self._eager(slot["sx"], slot["st"])
        except RuntimeError as e:
            text = f"{e} {e.__context__ or ''}"
            if "uncaptured work" in text or "StreamCaptureIsolation" in text:
                raise RuntimeError(
                    "TrainStep: the capture of the step's CUDA graph met work outside the capture -- most "
                    "likely the autograd graph of an eager forward through this model is still alive (a loss "
                    "or output tensor of it is referenced).  Delete it before the step captures (after "
                    f"{self.warmup} warm-up calls), or construct the step with graph=False.") from e
            raise
        if step_before is not None:
            opt._step = step_before  # capture does not execute; replays advance the counter
        self.launches_per_step = nv.launch_count()
        slot.update(graph=g, loss=loss, stats=stats)

    def _refresh_stale_packs(self):
        """A captured step re-packs the operand copies of the weights at its END, so weights changed
        from outside between two replays (load_state_dict, load_model, a manual copy_) would be
        multiplied from stale packs in the next forward.  torch bumps a tensor's version on every
        in-place write, which makes such a change cheap to notice; the packs are then rebuilt
        before the replay."""
        from .ops.modules import pack_weights, packed_layers
        if self._packed is None:
            self._packed = packed_layers(self.model)
        if any(m._packed_version != m.weight._version or m._packed_ptr != m.weight.data_ptr()
               for m in self._packed):
            pack_weights(self._packed)

    def __call__(self, x, t):
        if not self.use_graph:
            if x.device != self.device:
                x, t = x.to(self.device, non_blocking=True), t.to(self.device, non_blocking=True)
            loss, self._stats = self._eager(x, t)
            return loss
        slot = self._slot(x, t)
        if slot["graph"] is None:
            if slot["calls"] < self.warmup:
                slot["calls"] += 1
                loss, self._stats = self._eager(slot["sx"], slot["st"])
                return loss
            torch.cuda.synchronize(self.device)
            self._capture(slot)
        self._refresh_stale_packs()
        opt = self.optimizer
        if hasattr(opt, "refresh_hyper"):
            # the graph re-reads the pinned block; if the values changed (Adam: every step, SGD: when a
            # scheduler moved lr) refresh_hyper first waits for the previous replay's read of it
            opt.refresh_hyper(enqueue_copy=False)
        slot["graph"].replay()
        if hasattr(opt, "guard_hyper"):
            self._done = torch.cuda.Event()
            self._done.record()
            opt.guard_hyper(self._done)
        if hasattr(opt, "_step"):
            opt._step += 1
        self._stats = slot["stats"]
        return slot["loss"]

    @property
    def last_stats(self):
        """Device tensor [loss, #correct] of the last step (``CrossEntropyLoss`` criterion)."""
        return self._stats
