"""Single-process, multi-GPU data parallelism (``nn.DataParallel`` / ``MirroredStrategy`` style).

The reference's ``nn.DataParallel`` (pytorch/data_parallel.py:70-72) re-broadcasts all 92 MB of
parameters from GPU 0 at EVERY forward, scatters the input, runs one Python thread per GPU,
gathers logits and reduces all gradients onto GPU 0, which alone runs the optimizer
(SURVEY.md section 3.2).  This engine keeps the call contract -- input on GPU 0, logits on
GPU 0, ``.module`` is the original model, ``state_dict()`` keys carry ``module.`` -- but:

* replicas are persistent; the fused reducer updates every replica's parameters identically,
  so the per-step parameter broadcast (K-C5) does not exist;
* scatter is folded into each replica's first kernel, which reads its batch slice from GPU 0's
  input tensor over NVLink (K-C4); gather is folded into the last kernel, which writes fp32
  logits (or adds its share of the loss) straight into GPU 0's memory (K-C6);
* gradients are reduced by the same peer-memory fused all-reduce + optimizer kernels as the
  multi-process path (K-C7), one launch per bucket per GPU, instead of reduce-to-GPU-0;
* one thread drives all GPUs: :class:`DataParallelStep` replays ONE CUDA graph per GPU per step
  (forward, loss, backward, fused reduce+update, weight re-pack), which removes the GIL-bound
  thread-per-GPU structure that makes the reference's DataParallel its worst scaler.
"""
from __future__ import annotations

import copy
import gc
import weakref

import torch
import torch.nn as nn

from .. import _native as nv
from ..ops import functional as fn
from ..ops.modules import OutputLayout, pack_weights, packed_layers
from .arena import BufferArena, ParamArena
from .peer import PeerComm


class _GatherLogits(torch.autograd.Function):
    """[n_g,1,1,Cp] bf16 on GPU g  ->  rows of one [N,C] fp32 tensor on GPU 0.

    Forward: every replica's conversion kernel stores into GPU 0's buffer (peer write).
    Backward: every replica's conversion kernel loads its rows of dlogits from GPU 0 (peer read)
    and leaves the bf16 gradient on its own GPU, so replica backwards run on their own devices.
    """

    @staticmethod
    def forward(ctx, C, out_device, world, prefilled, *xs):
        ctx.world = world
        N = sum(x.shape[0] for x in xs)
        # ``prefilled``: the logits buffer on GPU 0 that every replica's last layer has already stored
        # its rows into from its GEMM epilogue (K-C6); otherwise a conversion kernel per replica does it
        out = prefilled if prefilled is not None else torch.empty((N, C), dtype=torch.float32, device=out_device)
        main = torch.cuda.current_stream(out_device)
        row, meta = 0, []
        for x in xs:
            n, Cp = x.shape[0], x.shape[-1]
            with torch.cuda.device(x.device):
                st = torch.cuda.current_stream(x.device)
                if prefilled is None:
                    nv.call("unpad_cast", nv.ptr(x), out[row: row + n].data_ptr(), n, C, Cp, fn.act_of(x), st.cuda_stream)
                ev = torch.cuda.Event()
                ev.record(st)
            main.wait_event(ev)
            meta.append((x.device, row, n, Cp, x.dtype))
            row += n
        ctx.meta, ctx.C = meta, C
        return out

    @staticmethod
    def backward(ctx, dy):
        # The user's loss is a mean over the GLOBAL batch, so a replica's parameter gradient is only
        # its shard's share of it; the bucket reduction then AVERAGES over replicas (the kernels'
        # 1/world).  Scaling the logits' gradient by world makes that average the true gradient --
        # the same correction DataParallelStep's shard loss applies (world / total).
        dy = (dy * float(ctx.world)).contiguous()
        ev0 = torch.cuda.Event()
        ev0.record(torch.cuda.current_stream(dy.device))
        grads = []
        for dev, row, n, Cp, dtype in ctx.meta:
            with torch.cuda.device(dev):
                st = torch.cuda.current_stream(dev)
                st.wait_event(ev0)
                dx = torch.empty((n, 1, 1, Cp), dtype=dtype, device=dev)
                nv.call("pad_cast", dy[row: row + n].data_ptr(), nv.ptr(dx), n, ctx.C, Cp, fn.act_of(dtype), st.cuda_stream)
            grads.append(dx)
        return (None, None, None, None) + tuple(grads)


class _ShardLoss(torch.autograd.Function):
    """Softmax cross entropy of one replica's shard of a global batch of ``total`` samples.

    The kernel adds this shard's share of the mean loss and its correct-prediction count into
    ``stats`` -- a 2-float buffer that may live on another GPU -- and produces dlogits scaled by
    world/total: the reducer then averages over the ``world`` replicas, which yields exactly the
    mean gradient over the global batch even when the shards are unequal.
    """

    @staticmethod
    def forward(ctx, logits, targets_ptr, stats_ptr, total, world):
        # targets / stats are raw device pointers (GPU 0 memory): handing GPU-0 tensors to
        # autograd would make the engine synchronise GPU 0's streams with this replica's
        logits = logits.contiguous()
        n, C = logits.shape
        dl = torch.empty_like(logits)
        nv.call("softmax_ce", nv.ptr(logits), targets_ptr, stats_ptr, 0, nv.ptr(dl), n, C, C,
                1.0 / total, float(world) / total, nv.stream_ptr())
        ctx.save_for_backward(dl)
        return torch.zeros((), device=logits.device)

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl, None, None, None, None


class DataParallel(nn.Module):
    """``model = DataParallel(model)`` then use it like ``nn.DataParallel(model)``."""

    def __init__(self, module, device_ids=None, bucket_cap_mb=25.0, first_bucket_mb=1.0):
        super().__init__()
        if not torch.cuda.is_available():
            raise RuntimeError("DataParallel needs CUDA devices; use wrap() with backend='gloo' on CPU")
        ids = list(range(torch.cuda.device_count())) if device_ids is None else list(device_ids)
        self.devices = [torch.device("cuda", i) for i in ids]
        self.world = len(self.devices)
        self.module = module.to(self.devices[0])
        # replicas live in a plain list: only replica 0 is a registered sub-module, so
        # parameters() / state_dict() describe ONE model, with the "module." prefix
        self.replicas = [self.module] + [copy.deepcopy(self.module).to(d) for d in self.devices[1:]]
        self.comms = PeerComm.create_local(self.devices)
        self.arenas, self.buffer_arenas, self.packed = [], [], []
        for rep, dev in zip(self.replicas, self.devices):
            with torch.cuda.device(dev):
                self.arenas.append(ParamArena(rep.parameters(), dev, peer=None, bucket_cap_mb=bucket_cap_mb,
                                              first_bucket_mb=first_bucket_mb))
                self.buffer_arenas.append(BufferArena(list(rep.buffers()), dev, peer=None))
                self.packed.append(packed_layers(rep))
        numel = self.arenas[0].numel
        pbufs = PeerComm.alloc_local(self.comms, numel * 4)
        gbufs = PeerComm.alloc_local(self.comms, numel * 4)
        bbufs = PeerComm.alloc_local(self.comms, self.buffer_arenas[0].nbytes)
        for g, dev in enumerate(self.devices):
            with torch.cuda.device(dev):
                self.arenas[g].peer = self.comms[g]
                self.arenas[g].adopt_symmetric(pbufs[g], gbufs[g])
                self.buffer_arenas[g].adopt_symmetric(bbufs[g])
            self.arenas[g].reducer = self      # lets FusedSGD(model.parameters()) find this engine
        self.optimizer = None
        self.state_sharded = False
        outs = [m for m in self.module.modules() if isinstance(m, OutputLayout)]
        if not outs:
            raise RuntimeError("DataParallel needs a model built on this package's op layer "
                               "(it must end in ops.OutputLayout)")
        self.out_features = outs[-1].features
        for d in self.devices:
            torch.cuda.synchronize(d)

    # ------------------------------------------------------------------ helpers
    @property
    def arena(self):
        return self.arenas[0]

    def shard_bounds(self, N):
        """Balanced split of a batch of N over the replicas (the first N % world shards get one
        more sample).  Every replica must own at least one sample: its collectives are part of every
        step, and a replica without a backward pass would contribute stale gradients."""
        if N < self.world:
            raise ValueError(f"batch of {N} samples cannot be split over {self.world} GPUs: "
                             "need at least one sample per replica")
        base, rem = divmod(N, self.world)
        bounds, lo = [], 0
        for g in range(self.world):
            hi = lo + base + (1 if g < rem else 0)
            bounds.append((lo, hi))
            lo = hi
        return bounds

    def sync_buffers_on(self, g):
        """Replica 0's BatchNorm running statistics become replica g's (K-C2)."""
        if self.world > 1 and self.buffer_arenas[0].nbytes > 16:
            self.comms[g].broadcast(self.buffer_arenas[g].buf, 0, self.buffer_arenas[g].nbytes, root=0,
                                    max_blocks=4)

    def run_replica_body(self, g, x_slice):
        """Replica g up to (not including) its OutputLayout; ``x_slice`` may be peer memory."""
        rep = self.replicas[g]
        last = [m for m in rep.modules() if isinstance(m, OutputLayout)][-1]
        last._bypass = True
        try:
            return rep(x_slice)
        finally:
            last._bypass = False

    def reduce_on(self, g, fused=True):
        """Replica g's launches of the bucket collectives (same bucket order on every GPU)."""
        opt, a, pc = self.optimizer, self.arenas[g], self.comms[g]
        if self.world == 1:
            if fused:
                opt._local_update_on(g)
            return
        # the buckets touch disjoint memory: one closing "done" exchange for all of them (see
        # csrc/comm/collectives.cu), not one per bucket
        pc.defer_done(True)
        try:
            for bucket in a.buckets:
                off_b, n = bucket.offset * 4, bucket.numel
                two = pc.choose_two_shot(n * 4)
                self.state_sharded = self.state_sharded or bool(two)
                if fused:
                    m, v = opt._state_on(g)
                    pc.reduce_update(a.grad_buf, a.param_buf, m, v, opt._hyper_on(g), off_b, n, opt.opt_code,
                                     two_shot=two, max_blocks=64)
                else:
                    pc.allreduce(a.grad_buf, off_b, n, scale=1.0 / self.world, max_blocks=64)
            pc.barrier()
        finally:
            pc.defer_done(False)

    # ------------------------------------------------------------------ nn.DataParallel contract
    def forward(self, x):
        """Input on GPU 0 (anything else is copied there), logits on GPU 0, autograd through all GPUs."""
        d0 = self.devices[0]
        if x.device != d0:
            x = x.to(d0, non_blocking=True)
        train = self.training and torch.is_grad_enabled()
        if train and self.optimizer is not None:
            self.optimizer.refresh_hyper()
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(d0))
        # Kernels that wait for another GPU are launched for ALL GPUs back to back, with
        # nothing in between that could block the host (an allocation, a synchronous copy):
        # a GPU spinning on a peer whose kernel the host has not launched yet would deadlock.
        for g, dev in enumerate(self.devices):
            with torch.cuda.device(dev):
                torch.cuda.current_stream(dev).wait_event(ev)
                if train:
                    self.sync_buffers_on(g)
        outs = []
        # gather folded into the replicas' last layer: its epilogue stores fp32 logits straight into this
        # buffer on GPU 0 (peer write over NVLink) when the model ends in Linear -> OutputLayout
        heads = self._gather_heads()
        logits = (torch.empty((x.shape[0], self.out_features), dtype=torch.float32, device=d0)
                  if heads is not None else None)
        for g, (lo, hi) in enumerate(self.shard_bounds(x.shape[0])):
            if lo >= hi:
                continue
            with torch.cuda.device(self.devices[g]):
                if heads is not None:
                    heads[g]._gather_dst = (logits[lo:hi].data_ptr(), self.out_features)
                try:
                    outs.append(self.run_replica_body(g, x[lo:hi]))
                finally:
                    if heads is not None:
                        heads[g]._gather_dst = None
        out = _GatherLogits.apply(self.out_features, d0, self.world, logits, *outs)
        # DataParallelStep refuses to capture while this call's autograd graph is still alive
        self._eager_graph = weakref.ref(out.grad_fn) if out.grad_fn is not None else None
        return out

    def eager_graph_alive(self):
        """True while the autograd graph of the last eager ``forward`` call is still referenced (by a
        loss, an output, a saved list ...), after a garbage collection has had its chance."""
        ref = getattr(self, "_eager_graph", None)
        if ref is None or ref() is None:
            return False
        gc.collect()
        return ref() is not None

    def _gather_heads(self):
        """The replicas' final Linear layers when the model ends in ``Linear -> OutputLayout`` with
        matching widths and no fused activation (then their epilogue can write GPU 0's logits);
        None otherwise."""
        if not hasattr(self, "_heads"):
            from ..ops.modules import Linear
            heads = []
            for rep in self.replicas:
                mods = [m for m in rep.modules() if len(list(m.children())) == 0]
                ok = (len(mods) >= 2 and isinstance(mods[-1], OutputLayout) and isinstance(mods[-2], Linear)
                      and mods[-2].out_features == self.out_features and not mods[-2].fuse_relu)
                heads.append(mods[-2] if ok else None)
            self._heads = heads if all(h is not None for h in heads) else None
        return self._heads

    # ------------------------------------------------------------------ reducer interface (FusedSGD)
    def finish_step(self, fused=True):
        """Called by ``optimizer.step()`` after backward: fused all-reduce+update per bucket per GPU."""
        for g, dev in enumerate(self.devices):
            with torch.cuda.device(dev):
                self.reduce_on(g, fused)

    def wait_gradients(self):
        """For stock torch optimizers: averages gradients over replicas (call before step())."""
        self.finish_step(fused=False)

    def repack_all(self, caches):
        for g, dev in enumerate(self.devices):
            with torch.cuda.device(dev):
                if self.packed[g]:
                    pack_weights(self.packed[g], caches[g])

    def train(self, mode=True):
        super().train(mode)
        for rep in self.replicas[1:]:
            rep.train(mode)
        return self

    def close(self):
        for pc in self.comms:
            pc.close()


class DataParallelStep:
    """One training step over all replicas: ``loss = step(x, y)`` (x, y on CPU pinned or GPU 0).

    After ``warmup`` eager iterations the per-replica work -- BN-stat sync, forward (first kernel
    reads its shard of the input from GPU 0), loss (adds its share into GPU 0's loss buffer),
    backward, fused bucket all-reduce + optimizer update, weight re-pack -- is captured into one
    CUDA graph per GPU and replayed.  Returns a device tensor ``[mean loss, #correct]`` on GPU 0.

    An autograd graph built through ``DataParallel.forward`` (the eager call contract) must not be
    alive when the capture happens: CUDA would report a dependency of the captured stream on
    uncaptured work.  The step checks for it and raises a RuntimeError that says so -- drop the
    eager call's ``loss`` / output first.
    """

    def __init__(self, model: DataParallel, optimizer, graph=True, warmup=2):
        self.model, self.opt = model, optimizer
        self.use_graph, self.warmup = graph, warmup
        self.d0 = model.devices[0]
        self._sx = self._st = None
        self._graphs = None
        self._calls = 0
        self._stats = torch.zeros(2, device=self.d0)
        self._streams = [torch.cuda.Stream(d) for d in model.devices]
        self._done = None
        self.launches_per_step = None

    def _replica_step(self, g, phase=None):
        """Replica g's whole step (``phase=None``, used under graph capture where nothing runs
        on the host side), or one phase of it: eager execution keeps the cross-GPU kernels of
        all replicas (phases 0 and 2) free of interleaved host work that might block."""
        m, opt = self.model, self.opt
        lo, hi = m.shard_bounds(self._sx.shape[0])[g]
        if lo >= hi:
            return
        if phase in (None, 0):
            m.sync_buffers_on(g)
        if phase in (None, 1):
            self._fwd_bwd(g, lo, hi)
        if phase in (None, 2):
            m.reduce_on(g, fused=True)
        if phase in (None, 3) and m.packed[g]:
            pack_weights(m.packed[g], opt._pack_caches[g])

    def _fwd_bwd(self, g, lo, hi):
        m, opt = self.model, self.opt
        opt._hyper_devs[g].copy_(opt._hyper_host, non_blocking=True)
        out = m.run_replica_body(g, self._sx[lo:hi])
        logits = fn.to_f32(out, m.out_features)
        loss = _ShardLoss.apply(logits, self._st[lo:hi].data_ptr(), self._stats.data_ptr(), self._sx.shape[0], m.world)
        for p in m.arenas[g].params:
            p._dtdl_fresh = True
        loss.backward()

    def _run_all(self, capture=False):
        m = self.model
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.d0))
        graphs = []
        for g, dev in enumerate(m.devices):
            with torch.cuda.device(dev):
                st = self._streams[g]
                st.wait_event(ev)
                if capture:
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr, stream=st, capture_error_mode="thread_local"):
                        self._replica_step(g)
                    graphs.append(gr)
                elif self._graphs is not None:
                    with torch.cuda.stream(st):
                        self._graphs[g].replay()
        if not capture and self._graphs is None:
            for phase in range(4):
                for g, dev in enumerate(m.devices):
                    with torch.cuda.device(dev), torch.cuda.stream(self._streams[g]):
                        self._replica_step(g, phase)
        return graphs

    def _join(self):
        main = torch.cuda.current_stream(self.d0)
        for g, dev in enumerate(self.model.devices):
            ev = torch.cuda.Event()
            with torch.cuda.device(dev):
                ev.record(self._streams[g])
            main.wait_event(ev)

    def __call__(self, x, t):
        m, opt = self.model, self.opt
        with torch.cuda.device(self.d0):
            if self._sx is None or self._sx.shape != x.shape:
                self._sx = torch.empty(x.shape, dtype=torch.float32, device=self.d0)
                self._st = torch.empty(t.shape, dtype=torch.int64, device=self.d0)
                self._graphs, self._calls = None, 0
            opt.refresh_hyper(enqueue_copy=False)   # waits for the previous step's read if values changed
            self._sx.copy_(x, non_blocking=True)
            self._st.copy_(t, non_blocking=True)
            self._stats.zero_()
        if self.use_graph and self._graphs is None and self._calls >= self.warmup:
            if m.eager_graph_alive():
                raise RuntimeError(
                    "DataParallelStep: the autograd graph of an earlier eager model(x) call is still alive "
                    "(a loss or output tensor of it is referenced); delete it before the step captures its "
                    "CUDA graphs, or construct the step with graph=False")
            for d in m.devices:
                torch.cuda.synchronize(d)
            nv.reset_launch_count()
            graphs = self._run_all(capture=True)
            self.launches_per_step = nv.launch_count()
            self._graphs = graphs
        self._calls += 1
        self._run_all()
        self._join()
        opt._step += 1
        self._done = torch.cuda.Event()
        self._done.record(torch.cuda.current_stream(self.d0))
        opt.guard_hyper(self._done)
        with torch.cuda.device(self.d0):
            return self._stats.clone()
