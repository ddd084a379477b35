"""``wrap(model)``: the drop-in data-parallel wrapper (one process per GPU).

Same contract as ``torch.nn.parallel.DistributedDataParallel`` as the reference uses it
(pytorch/distributed_data_parallel.py:74,128-133): the wrapped model is called like the bare
one, ``.module`` is the original, ``state_dict()`` keys carry the ``module.`` prefix
(pytorch/data_parallel.py:87-90) while ``.module.state_dict()`` is bare (:108-114), parameters
and buffers are broadcast from rank 0 at construction, buffers (BatchNorm running statistics)
are re-broadcast at every forward, and gradients are averaged over ranks during backward,
bucket by bucket, overlapped with the rest of backward on a side stream.

What is different underneath (``backend="b200"``): parameters, gradients and buffers live in
symmetric peer-memory arenas; a bucket's all-reduce is a hand-written kernel reading /
writing peer memory over NVLink, and when a fused optimizer is attached the same kernel applies
1/W, weight decay and the SGD/Adam update and scatters updated parameters -- so there is no
separate ``optimizer.step()`` pass.  ``backend="gloo"`` (CPU) and ``"nccl"`` run the same
bucket logic over ``torch.distributed`` collectives.
"""
from __future__ import annotations

import os
import time

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import functional as fn
from ..ops.modules import packed_layers
from . import rendezvous
from .arena import BufferArena, ParamArena

_NVTX = os.environ.get("DTDL_NVTX") == "1"  # NVTX ranges around bucket launches (Nsight timelines)


class Reducer:
    """Bucketed gradient reducer.  Buckets launch strictly in index order on every rank, so
    all ranks issue the same sequence of collectives whatever the local gradient order."""

    def __init__(self, module, comm, bucket_cap_mb=25.0, first_bucket_mb=1.0, broadcast_buffers=True,
                 grad_dtype="fp32"):
        # tuning overrides for measurements (tools/sweep_2gpu.sh): bucket sizes, buffer broadcast
        bucket_cap_mb = float(os.environ.get("DTDL_BUCKET_MB", bucket_cap_mb))
        first_bucket_mb = float(os.environ.get("DTDL_FIRST_BUCKET_MB", first_bucket_mb))
        if os.environ.get("DTDL_BROADCAST_BUFFERS") == "0":
            broadcast_buffers = False
        self.module, self.comm = module, comm
        self.world, self.rank = comm.size, comm.rank
        self.device = next(module.parameters()).device
        self.backend = comm.backend if self.world > 1 else ("b200" if self.device.type == "cuda" else "gloo")
        self.peer = comm.peer if (comm.backend == "b200" and self.device.type == "cuda") else None
        self.arena = ParamArena(module.parameters(), self.device, peer=self.peer,
                                bucket_cap_mb=bucket_cap_mb, first_bucket_mb=first_bucket_mb)
        self.arena.reducer = self
        self.buffers = BufferArena(list(module.buffers()), self.device, peer=self.peer)
        self.broadcast_buffers = broadcast_buffers
        self.packed_modules = packed_layers(module) if self.device.type == "cuda" else []
        # conv / linear layers whose weight lives in bucket b: their operand packs are rebuilt on the
        # communication stream right behind b's update, under the rest of backward (packs_in_buckets
        # tells the optimizer that step() has nothing left to re-pack)
        self.packs_in_buckets = self.device.type == "cuda"
        self._bucket_packed = {b.index: [] for b in self.arena.buckets}
        for m in self.packed_modules:
            b = self.arena.bucket_of.get(id(m.weight))
            if b is not None and getattr(m, "pack_src", None) is None:
                self._bucket_packed[b].append(m)
        self._bucket_pack_cache = {b.index: {} for b in self.arena.buckets}
        self._deferred = False
        self._updated_buckets = 0
        self.packs_current = False     # True after a step whose buckets all re-packed their layers
        # gradient transport type of the fused reducer: "fp32" (what DDP moves) or "bf16" (cast per bucket,
        # half the NVLink bytes, fp32 accumulation and update)
        if grad_dtype not in ("fp32", "bf16"):
            raise ValueError("grad_dtype must be 'fp32' or 'bf16'")
        self.grad_bf16 = grad_dtype == "bf16" and self.peer is not None and self.world > 1
        if self.grad_bf16:
            self.arena.enable_bf16_grads()
        self.optimizer = None
        self.state_sharded = False
        self.require_sync = True
        self.comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._pending = None
        self._next = 0
        self._works = []
        self._launched_any = False
        self._callback_queued = False
        # per-bucket timing of the collectives (eager steps only; DTDL_BUCKET_TIMING=1 or
        # reducer.bucket_timing = True), read back with bucket_report()
        self.bucket_timing = os.environ.get("DTDL_BUCKET_TIMING") == "1"
        self._bucket_events = []
        self._hook = fn.add_ready_hook(self._on_ready)
        for p in self.arena.params:
            p.register_post_accumulate_grad_hook(self._torch_hook)
        self.sync_parameters()

    # ------------------------------------------------------------------ construction-time sync
    def sync_parameters(self):
        """Rank 0's parameters and buffers become everyone's (K-C1)."""
        if self.world == 1:
            return
        if self.peer is not None:
            torch.cuda.synchronize(self.device)
            self.peer.broadcast(self.arena.param_buf, 0, self.arena.numel * 4, root=0)
            self.peer.broadcast(self.buffers.buf, 0, self.buffers.nbytes, root=0)
            torch.cuda.synchronize(self.device)
        else:
            dist.broadcast(self.arena.param_flat, src=0, group=self.comm.group)
            dist.broadcast(self.buffers.bytes, src=0, group=self.comm.group)
        for m in self.packed_modules:
            m._packed_version = None

    def sync_buffers(self):
        """Per-step BatchNorm running-stat sync from rank 0 (K-C2).

        torch.distributed back ends: at the start of every forward, as DDP does.  Peer back end: the
        broadcast rides on the communication stream in front of the step's first bucket (see
        ``_launch_impl``), i.e. it happens once this step's forward has updated rank 0's statistics
        and is overlapped with backward; every rank therefore starts the next forward from rank 0's
        values -- the state DDP's broadcast-before-forward produces -- without a cross-GPU wait on the
        critical path of the forward pass."""
        if self.world == 1 or not self.broadcast_buffers or self.buffers.nbytes <= 16:
            return
        if self.peer is None:
            dist.broadcast(self.buffers.bytes, src=0, group=self.comm.group)

    def _sync_buffers_on_comm_stream(self):
        if self.world == 1 or not self.broadcast_buffers or self.buffers.nbytes <= 16:
            return
        self.peer.broadcast(self.buffers.buf, 0, self.buffers.nbytes, root=0, max_blocks=4,
                            stream=self.comm_stream)

    # ------------------------------------------------------------------ backward-time reduction
    def prepare(self):
        """Start of an iteration: every bucket waits for all of its gradients."""
        if self._pending is not None and self._next not in (0, len(self.arena.buckets)):
            self.flush()
        self._pending = [len(b.params) for b in self.arena.buckets]
        self._seen = set()
        # the stream the user's step runs on: end-of-backward joins are applied to THIS stream (the
        # autograd callback may run on an engine thread whose current stream is a different one)
        self._user_stream = torch.cuda.current_stream(self.device) if self.device.type == "cuda" else None
        self._next = 0
        self._launched_any = False
        self._updated_buckets = 0
        self.packs_current = False
        if self._deferred and self.peer is not None:
            # a previous iteration was abandoned between its first bucket and finish_step (an exception,
            # a backward without a step): close its sequence so that later collectives are self-contained
            self.peer.barrier(stream=self.comm_stream)
            self.peer.defer_done(False)
            self._deferred = False
        if self.optimizer is not None:
            self.optimizer.refresh_hyper()

    def _torch_hook(self, p):
        # gradients produced by stock torch ops land in p.grad; move them into the arena
        # (autograd accumulates into p.grad itself, so p.grad always holds the full value)
        g = p._dtdl_grad
        if p.grad is not None and p.grad.data_ptr() != g.data_ptr():
            g.copy_(p.grad)
            p.grad = g
        p._dtdl_fresh = False
        self._on_ready(p)

    def _on_ready(self, p):
        if self._pending is None or getattr(p, "_dtdl_arena", None) is not self.arena:
            return
        if not self._callback_queued:
            # end-of-backward: launch leftovers; without a fused optimizer also join, so that
            # p.grad holds the averaged gradient when backward() returns (DDP's contract)
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._backward_done)
        # A parameter can be announced twice in one backward (by the kernel that wrote its
        # gradient AND by autograd's post-accumulate hook, which torch also fires for parameters
        # whose Function returned no gradient): count it once.
        if id(p) in self._seen:
            return
        self._seen.add(id(p))
        b = self.arena.bucket_of[id(p)]
        self._pending[b] -= 1
        while self._next < len(self.arena.buckets) and self._pending[self._next] <= 0:
            self._launch(self.arena.buckets[self._next])
            self._next += 1

    def _backward_done(self):
        self._callback_queued = False
        fn.join_side_streams()   # idempotent: the weight-gradient branch joins the main chain first
        if self.peer is not None and self.peer._trace is not None:
            self.peer.trace_mark(stream=getattr(self, "_user_stream", None))   # "backward has ended"
        if not self.require_sync:
            return
        if self.optimizer is None:
            self.finish_step(fused=False)
        else:
            self.flush()

    def _fused(self):
        return self.optimizer is not None and self.require_sync

    def _launch(self, bucket):
        if not self.require_sync:
            return
        if _NVTX:
            torch.cuda.nvtx.range_push(f"dtdl.bucket{bucket.index}.{bucket.numel * 4 >> 10}KiB")
        try:
            self._launch_impl(bucket)
        finally:
            if _NVTX:
                torch.cuda.nvtx.range_pop()

    def _join_producers(self):
        """The communication stream waits for everything that may have produced this bucket's
        gradients: the stream this hook runs on, the user's stream, and the side stream that carries
        weight gradients during graph capture."""
        streams = [torch.cuda.current_stream(self.device)]
        for extra in (getattr(self, "_user_stream", None), fn.forked_side_stream(self.device)):
            if extra is not None and all(extra != s for s in streams):
                streams.append(extra)
        for s in streams:
            ev = torch.cuda.Event()
            ev.record(s)
            self.comm_stream.wait_event(ev)

    def _repack_bucket(self, bucket):
        mods = self._bucket_packed.get(bucket.index)
        if mods:
            from ..ops.modules import pack_weights
            with torch.cuda.stream(self.comm_stream):
                pack_weights(mods, self._bucket_pack_cache[bucket.index])

    def _launch_impl(self, bucket):
        off_b, n = bucket.offset * 4, bucket.numel
        a = self.arena
        if self.peer is not None and self.world > 1:
            self._join_producers()
            if not self._launched_any:
                # first bucket of the step: from here to finish_step the collectives run back to back
                # on the communication stream and share ONE closing "done" exchange
                self.peer.defer_done(True)
                self._deferred = True
                self._sync_buffers_on_comm_stream()
            timed = self.bucket_timing and not torch.cuda.is_current_stream_capturing()
            if timed:
                e0 = torch.cuda.Event(enable_timing=True)
                e0.record(self.comm_stream)
            two = self.peer.choose_two_shot(n * 4)
            if self._fused():
                opt = self.optimizer
                m = opt._state_flat[opt.state_names[0]]
                v = opt._state_flat[opt.state_names[1]] if len(opt.state_names) > 1 else None
                if self.grad_bf16:
                    from .. import _native as nv
                    two = two or 1       # the bf16 reducers shard like two-shot / multicast
                    nv.call("cast_f32_bf16", a.grad_flat[bucket.offset:].data_ptr(),
                            a.grad16_flat[bucket.offset:].data_ptr(), n, self.comm_stream.cuda_stream)
                    self.state_sharded = True
                    self.peer.reduce_update16(a.grad16_buf, a.param_buf, m, v, opt._hyper_dev, off_b, n,
                                              opt.opt_code, stream=self.comm_stream)
                else:
                    self.state_sharded = self.state_sharded or bool(two)
                    self.peer.reduce_update(a.grad_buf, a.param_buf, m, v, opt._hyper_dev, off_b, n,
                                            opt.opt_code, two_shot=two, stream=self.comm_stream)
                self._repack_bucket(bucket)
            else:
                self.peer.allreduce(a.grad_buf, off_b, n, scale=1.0 / self.world, stream=self.comm_stream)
            if timed:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record(self.comm_stream)
                # an in-place plain all-reduce never runs the one-shot schedule (peer.allreduce)
                variant = "nvls" if two == 2 else ("two-shot" if (two == 1 or not self._fused()) else "one-shot")
                kind = ("fused " if self._fused() else "") + variant
                self._bucket_events.append((bucket, kind, e0, e1))
        elif self.world > 1:
            g = a.grad_flat[bucket.offset: bucket.offset + n]
            t0 = time.perf_counter() if self.bucket_timing else None
            work = dist.all_reduce(g, group=self.comm.group, async_op=True)
            self._works.append((work, bucket, t0))
        elif self._fused() and self.device.type == "cuda":
            # one GPU: the bucket's optimizer update and re-pack run on the side stream as soon as
            # its gradients are complete, under the rest of backward, instead of after it
            self._join_producers()
            opt = self.optimizer
            m = opt._state_flat[opt.state_names[0]]
            v = opt._state_flat[opt.state_names[1]] if len(opt.state_names) > 1 else None
            sl = slice(bucket.offset, bucket.offset + n)
            from .. import _native as nv
            nv.call("local_update", nv.ptr(a.param_flat[sl]), nv.ptr(a.grad_flat[sl]), nv.ptr(m[sl]),
                    nv.ptr(v[sl]) if v is not None else 0, nv.ptr(opt._hyper_dev), n, opt.opt_code,
                    self.comm_stream.cuda_stream)
            self._repack_bucket(bucket)
            self._updated_buckets += 1
        self._launched_any = True

    def flush(self):
        """Launches whatever has not been launched (parameters that received no gradient)."""
        if self._pending is None:
            return
        while self._next < len(self.arena.buckets):
            self._launch(self.arena.buckets[self._next])
            self._next += 1

    def finish_step(self, fused):
        """Called by ``optimizer.step()`` (fused) or by the wrapper before a torch optimizer."""
        self.flush()
        a = self.arena
        if self.peer is not None and self.world > 1:
            if self._deferred:
                # the step's single "done" exchange: no rank reads or writes our arenas after this
                self.peer.barrier(stream=self.comm_stream)
                self.peer.defer_done(False)
                self._deferred = False
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
            if getattr(self, "_user_stream", None) is not None:
                self._user_stream.wait_stream(self.comm_stream)
            self.packs_current = bool(fused and self.require_sync and self.packs_in_buckets)
        elif self.world > 1:
            for work, bucket, t0 in self._works:
                work.wait()
                if t0 is not None:   # host clock from launch to completion: an upper bound
                    self._bucket_events.append((bucket, self.backend, (time.perf_counter() - t0) * 1e3, None))
                g = a.grad_flat[bucket.offset: bucket.offset + bucket.numel]
                g.div_(self.world)
                if fused:
                    self.optimizer._update_torch(a.grad_flat, bucket.offset, bucket.numel)
            self._works = []
        if fused and self.world == 1:
            if self._updated_buckets == len(a.buckets) and self.comm_stream is not None:
                torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
                if getattr(self, "_user_stream", None) is not None:
                    self._user_stream.wait_stream(self.comm_stream)
                self.packs_current = True
            else:
                assert self._updated_buckets == 0, "a step updated only some of its buckets"
                self.optimizer._local_update()
                self.packs_current = False
        self._pending = None

    def wait(self):
        """Makes reduced gradients visible to the current stream (non-fused use)."""
        self.finish_step(fused=False)

    def close(self):
        """Detaches this reducer from the op layer's gradient-ready notifications (a process that
        wraps several models one after the other calls this before dropping a wrapped model)."""
        fn.remove_ready_hook(self._hook)
        self._pending = None

    def bucket_report(self, clear=True):
        """Per-bucket time and achieved bus bandwidth of the collectives timed since the last call
        (SURVEY.md section 5.5).  Bus bandwidth = 2 (W-1)/W * bytes / time, the figure NCCL's
        tests quote.  Peer back end: device time of the kernel on the communication stream
        (includes waiting for the slowest rank); torch.distributed back ends: host clock."""
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        rows = []
        for bucket, kind, a0, a1 in self._bucket_events:
            ms = a0.elapsed_time(a1) if a1 is not None else a0
            nbytes = bucket.numel * 4
            bus = 2.0 * (self.world - 1) / self.world * nbytes / (ms * 1e-3) / 1e9 if ms > 0 else float("inf")
            rows.append({"bucket": bucket.index, "bytes": nbytes, "ms": ms, "bus_GBps": bus, "kind": kind,
                         "world": self.world})
        if clear:
            self._bucket_events = []
        return rows


class DistributedModel(nn.Module):
    def __init__(self, module, comm=None, optimizer=None, bucket_cap_mb=25.0, first_bucket_mb=1.0,
                 broadcast_buffers=True, device_ids=None, output_device=None, dtype=None, grad_dtype="fp32"):
        super().__init__()
        if dtype is not None:
            from ..ops.modules import set_precision
            set_precision(module, dtype)
        comm = comm or rendezvous.current() or rendezvous.init()
        self.module = module
        self.comm = comm
        self.reducer = Reducer(module, comm, bucket_cap_mb, first_bucket_mb, broadcast_buffers, grad_dtype)
        if optimizer is not None:
            optimizer.attach(self.reducer)

    def forward(self, *args, **kwargs):
        red = self.reducer
        if self.training and torch.is_grad_enabled():
            red.sync_buffers()
            red.prepare()
        return self.module(*args, **kwargs)

    # gradient accumulation: skip the collective, let gradients add up locally
    def no_sync(self):
        model = self

        class _NoSync:
            def __enter__(self_inner):
                model.reducer.require_sync = False

            def __exit__(self_inner, *exc):
                model.reducer.require_sync = True

        return _NoSync()

    def wait_gradients(self):
        """For stock torch optimizers: call before ``optimizer.step()`` (a fused optimizer does it)."""
        self.reducer.wait()

    def sync_parameters(self):
        """Rank 0's parameters and buffers become everyone's again (the broadcast ``wrap`` does at
        construction): call it on every rank after loading weights on rank 0 only."""
        self.reducer.sync_parameters()


def wrap(model, optimizer=None, comm=None, **kw) -> DistributedModel:
    """``net = wrap(net)`` replaces ``DistributedDataParallel(net, device_ids=[gpu])``.

    ``dtype="bf16"`` (default of the layers) or ``dtype="tf32"``: the activation / operand precision
    of the wrapped model's layers, see ``ops.set_precision``.  ``grad_dtype="bf16"``: gradients cross
    NVLink as bf16 (cast per bucket, fp32 accumulation and update); default ``"fp32"`` as in DDP."""
    return DistributedModel(model, comm=comm, optimizer=optimizer, **kw)


# torch spelling kept for drop-in use
DistributedDataParallel = DistributedModel
