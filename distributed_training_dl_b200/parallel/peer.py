"""Peer-memory runtime: symmetric buffers + the hand-written collectives over them.

One :class:`PeerComm` per GPU.  Two ways to build a group:

* ``PeerComm.create_multiprocess(comm)`` -- one process per GPU (the DDP launch style,
  pytorch/distributed_data_parallel.py:53-64).  Buffers are ``cudaMalloc``'ed and shared with
  CUDA IPC handles exchanged through the rendezvous store.
* ``PeerComm.create_local(devices)`` -- one process driving several GPUs (the
  ``nn.DataParallel`` / ``MirroredStrategy`` style, pytorch/data_parallel.py:70-72).  Peer
  access is enabled between the devices and raw pointers are shared directly.

A symmetric buffer is a :class:`SymmBuffer`: the local memory as a torch tensor plus the
table of every rank's base pointer that the kernels dereference over NVLink.

NVLS (``DTDL_NVLS=1`` or ``PeerComm.use_nvls = True`` before the first ``alloc``): buffers come
from the cuMem allocator of csrc/comm/vmm.cpp instead, are additionally bound to an NVSwitch
multicast object, and all-reduce / fused reduce+update / broadcast go through the multicast
address (``multimem.ld_reduce`` / ``multimem.st``).  Handles cross processes as POSIX file
descriptors over Unix sockets (:class:`FdChannel`).
"""
from __future__ import annotations

import ctypes
import os
import shutil
import socket
import struct
import tempfile

import torch

from .. import _native as nv


class _RawCudaMemory:
    """Exposes a raw device allocation to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.ptr, self.nbytes, self.owner = ptr, nbytes, owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None}


def tensor_from_ptr(ptr: int, nbytes: int, device: torch.device, owner=None) -> torch.Tensor:
    raw = _RawCudaMemory(ptr, nbytes, owner)
    t = torch.as_tensor(raw, device=device)
    if t.data_ptr() != ptr or t.device != torch.device(device):
        # torch resolved the pointer to another device and made a copy: the tensor would not
        # alias the symmetric memory any more
        raise RuntimeError(f"device pointer {ptr:#x} does not resolve to {device} (torch saw {t.device})")
    t._dtdl_raw = raw  # keep the holder (and thereby the allocation owner) alive
    return t


class FdChannel:
    """All-gather of file descriptors between the processes of one node (SCM_RIGHTS).

    Every rank listens on a Unix socket whose path it publishes in the rendezvous store; an
    exchange is "send my fds to every peer, then accept one message from every peer".  A sent
    message waits in the listener's queue until it is accepted, so no rank blocks another.
    """

    def __init__(self, rank: int, world: int, store, barrier):
        self.rank, self.world, self._barrier = rank, world, barrier
        self._dir = tempfile.mkdtemp(prefix="dtdl_fd_")
        path = os.path.join(self._dir, f"rank{rank}.sock")
        self._sock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self._sock.bind(path)
        self._sock.listen(4 * world)
        store.set(f"dtdl/fdsock/{rank}", path)
        self._paths = [store.get(f"dtdl/fdsock/{r}").decode() for r in range(world)]
        self._seq = 0
        barrier()

    def allgather(self, fds) -> list:
        """Every rank contributes a (possibly empty) list of fds; returns the list per rank.
        Received descriptors are new numbers in this process and are the caller's to close."""
        self._seq += 1
        out = [None] * self.world
        out[self.rank] = list(fds)
        conns = []
        for r in range(self.world):
            if r == self.rank:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(self._paths[r])
            socket.send_fds(c, [struct.pack("iii", self.rank, self._seq, len(fds))], list(fds))
            conns.append(c)
        for _ in range(self.world - 1):
            c, _addr = self._sock.accept()
            msg, rfds, _flags, _a = socket.recv_fds(c, 12, 16)
            src, seq, n = struct.unpack("iii", msg)
            if seq != self._seq or n != len(rfds):
                raise RuntimeError(f"fd exchange out of step: rank {src} sent seq {seq} ({len(rfds)}/{n} fds), "
                                   f"expected seq {self._seq}")
            out[src] = list(rfds)
            c.close()
        self._barrier()       # nobody starts the next exchange before this one is drained
        for c in conns:
            c.close()
        return out

    def close(self):
        self._sock.close()
        shutil.rmtree(self._dir, ignore_errors=True)


class SymmBuffer:
    """Same-sized allocation on every rank + the per-rank base pointer table."""

    def __init__(self, comm: "PeerComm", nbytes: int, local_ptr: int, peer_ptrs, opened, mc_ptr: int = 0,
                 vmm=None):
        self.comm, self.nbytes, self.local_ptr = comm, nbytes, local_ptr
        self.peer_ptrs = list(peer_ptrs)
        self._opened = opened  # IPC mappings to close
        self.mc_ptr = mc_ptr   # NVLS multicast address of the same bytes on every rank (0: none)
        self._vmm = vmm        # (mapped size, [mapped pointers], [handles]) of a cuMem-backed buffer
        self.ptr_array = (ctypes.c_void_p * len(self.peer_ptrs))(*self.peer_ptrs)
        self.bytes = tensor_from_ptr(local_ptr, nbytes, comm.device, owner=self)
        self._freed = False

    def view(self, dtype, offset_bytes=0, numel=None) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        numel = (self.nbytes - offset_bytes) // esz if numel is None else numel
        return self.bytes[offset_bytes: offset_bytes + numel * esz].view(dtype)

    def free(self):
        if self._freed:
            return
        self._freed = True
        if self._vmm is not None:
            size, ptrs, handles = self._vmm
            for p in ptrs:
                nv.lib().dtdl_vmm_unmap(ctypes.c_void_p(p), size)
            for h in handles:
                nv.lib().dtdl_vmm_release(h)
            return
        with torch.cuda.device(self.comm.device):
            for p in self._opened:
                nv.lib().dtdl_ipc_close_handle(p)
            nv.lib().dtdl_symm_free(self.local_ptr)


class PeerComm:
    def __init__(self, rank: int, world: int, device: torch.device):
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self.handle = ctypes.c_void_p()
        self.flags_ptr = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            nv.call("comm_create", rank, world, ctypes.byref(self.handle), ctypes.byref(self.flags_ptr))
        self._store = None          # torch.distributed store (multi-process mode)
        self._local_group = None    # list of sibling PeerComm (single-process mode)
        self._seq = 0
        self._buffers = []
        # SM budget of a collective kernel (leaves SMs to backward)
        self.max_blocks = int(os.environ.get("DTDL_COMM_BLOCKS", "32"))
        self.two_shot_min_bytes = 256 * 1024
        # NVLS: allocate through cuMem + multicast and reduce inside NVSwitch.  DTDL_NVLS = 1 / 0
        # forces it on / off; "auto" (default) turns it on from NVLS_AUTO_MIN_WORLD ranks up when
        # every rank can join a multicast team (decided in create_multiprocess / create_local).
        self.use_nvls = False
        self.nvls_min_bytes = int(os.environ.get("DTDL_NVLS_MIN_BYTES", "0"))
        self._fdchan = None
        self._deferred = False
        self._trace = None
        self._last_stream = None    # stream of the last collective (see _stream)

    # ------------------------------------------------------------------ construction
    @classmethod
    def create_multiprocess(cls, comm) -> "PeerComm":
        import torch.distributed as dist
        pc = cls(comm.rank, comm.size, comm.device)
        if comm.size > 1:
            pc._store = dist.distributed_c10d._get_default_store()
            pc._control = comm.control
            flags = pc._exchange_ipc(int(pc.flags_ptr.value), "flags")
            for r, p in enumerate(flags):
                if r != pc.rank:
                    nv.call("comm_set_peer_flags", pc.handle, r, p)
            oks = [None] * comm.size
            dist.all_gather_object(oks, pc._nvls_probe(), group=comm.control)
            pc._decide_nvls(all(oks))
        return pc

    @classmethod
    def create_local(cls, devices) -> list:
        devices = [torch.device(d) for d in devices]
        world = len(devices)
        for a in devices:
            for b in devices:
                if a != b:
                    nv.call("enable_peer_access", a.index, b.index)
        comms = [cls(r, world, d) for r, d in enumerate(devices)]
        for c in comms:
            c._local_group = comms
            for r, o in enumerate(comms):
                if r != c.rank:
                    nv.call("comm_set_peer_flags", c.handle, r, o.flags_ptr)
        ok = world > 1 and all(c._nvls_probe() for c in comms)
        for c in comms:
            c._decide_nvls(ok)
        return comms

    # With 2 ranks the multicast schedule moves 1.5x the bytes of two-shot over NVLink (a rank's own
    # copy also travels to the switch and back); from 4 ranks up it moves fewer (1 + 1/W vs
    # 2 (W-1)/W per direction) -- measured in profiles/allreduce_sweep_nvls_{2,8}gpu.md.
    NVLS_AUTO_MIN_WORLD = 4

    def _nvls_probe(self) -> bool:
        """Can this rank take part in multicast collectives?  Rank 0 also creates (and drops) a
        small multicast object: that is the call that fails when the fabric is not set up."""
        if os.environ.get("DTDL_NVLS", "auto") == "0" or self.world < 2 or not self.multicast_supported():
            return False
        dev = self.device.index or 0
        gran = nv.lib().dtdl_vmm_granularity(dev, self.world)
        if gran <= 0:
            return False
        if self.rank == 0:
            mc = ctypes.c_longlong()
            shareable = 0 if self._local_group is not None else 1
            if nv.lib().dtdl_mc_create(self.world, gran, shareable, ctypes.byref(mc)) != 0:
                return False
            nv.lib().dtdl_vmm_release(mc)
        return True

    def _decide_nvls(self, all_ok: bool):
        mode = os.environ.get("DTDL_NVLS", "auto")
        if mode == "1" and not all_ok and self.world > 1:
            raise nv.NativeError("DTDL_NVLS=1 but a rank cannot join an NVLS multicast team "
                                 "(no NVSwitch multicast support, or the fabric manager is not running)")
        self.use_nvls = bool(all_ok and (mode == "1" or (mode == "auto" and self.world >= self.NVLS_AUTO_MIN_WORLD)))

    # ------------------------------------------------------------------ IPC plumbing
    def _exchange_ipc(self, local_ptr: int, tag: str):
        """All ranks publish an IPC handle for ``local_ptr``; returns every rank's mapped pointer."""
        import torch.distributed as dist
        self._seq += 1
        key = f"dtdl/ipc/{tag}/{self._seq}"
        buf = ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            nv.call("ipc_get_handle", ctypes.c_void_p(local_ptr), buf)
        self._store.set(f"{key}/{self.rank}", buf.raw)
        ptrs = []
        for r in range(self.world):
            if r == self.rank:
                ptrs.append(local_ptr)
                continue
            raw = self._store.get(f"{key}/{r}")
            out = ctypes.c_void_p()
            with torch.cuda.device(self.device):
                nv.call("ipc_open_handle", ctypes.create_string_buffer(bytes(raw), 64), ctypes.byref(out))
            ptrs.append(int(out.value))
        dist.barrier(group=self._control)
        return ptrs

    # ------------------------------------------------------------------ cuMem / multicast plumbing
    def _vmm_size(self, nbytes: int):
        gran = nv.lib().dtdl_vmm_granularity(self.device.index or 0, self.world)
        if gran <= 0:
            raise nv.NativeError("cuMem / multicast granularity query failed: NVLS is not available here")
        return (nbytes + gran - 1) // gran * gran, gran

    def _alloc_vmm(self, nbytes: int) -> SymmBuffer:
        """cuMem-backed symmetric buffer with peer mappings and a multicast mapping (multi-process)."""
        import torch.distributed as dist
        dev = self.device.index or 0
        size, gran = self._vmm_size(nbytes)
        if self._fdchan is None:
            self._fdchan = FdChannel(self.rank, self.world, self._store,
                                     lambda: dist.barrier(group=self._control))
        ll, ci = ctypes.c_longlong, ctypes.c_int
        mem = ll()
        nv.call("vmm_create", dev, size, 1, ctypes.byref(mem))
        fd = ci()
        nv.call("vmm_export_fd", mem, ctypes.byref(fd))
        mine = [fd.value]
        mc = ll()
        if self.rank == 0:
            nv.call("mc_create", self.world, size, 1, ctypes.byref(mc))
            nv.call("vmm_export_fd", mc, ctypes.byref(fd))
            mine.append(fd.value)
        gathered = self._fdchan.allgather(mine)
        devs = (ci * 1)(dev)
        ptrs, mapped = [], []
        for r in range(self.world):
            h = mem
            if r != self.rank:
                h = ll()
                nv.call("vmm_import_fd", gathered[r][0], ctypes.byref(h))
            out = ctypes.c_void_p()
            nv.call("vmm_map", h, size, gran, devs, 1, ctypes.byref(out))
            if r != self.rank:
                nv.call("vmm_release", h)      # the mapping keeps the peer's memory alive
            ptrs.append(int(out.value))
            mapped.append(int(out.value))
        if self.rank != 0:
            nv.call("vmm_import_fd", gathered[0][1], ctypes.byref(mc))
        nv.call("mc_add_device", mc, dev)
        dist.barrier(group=self._control)      # binding needs the whole team added
        nv.call("mc_bind", mc, mem, size)
        out = ctypes.c_void_p()
        nv.call("vmm_map", mc, size, gran, devs, 1, ctypes.byref(out))
        mapped.append(int(out.value))
        for fds in gathered:
            for f in fds:
                os.close(f)
        buf = SymmBuffer(self, nbytes, ptrs[self.rank], ptrs, [], mc_ptr=int(out.value),
                         vmm=(size, mapped, [mem, mc]))
        buf.bytes.zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self._control)
        return buf

    def alloc(self, nbytes: int, tag: str = "buf") -> SymmBuffer:
        """Collective: every rank must call with the same size in the same order."""
        nbytes = (int(nbytes) + 255) // 256 * 256
        if self._local_group is not None:
            raise RuntimeError("use PeerComm.alloc_local(comms, nbytes) in single-process mode")
        if self.use_nvls and self.world > 1:
            buf = self._alloc_vmm(nbytes)
            self._buffers.append(buf)
            return buf
        out = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            nv.call("symm_alloc", nbytes, ctypes.byref(out))
        local = int(out.value)
        if self.world > 1:
            ptrs = self._exchange_ipc(local, tag)
            opened = [p for r, p in enumerate(ptrs) if r != self.rank]
        else:
            ptrs, opened = [local], []
        buf = SymmBuffer(self, nbytes, local, ptrs, opened)
        self._buffers.append(buf)
        return buf

    @staticmethod
    def _alloc_local_vmm(comms, nbytes: int):
        """Single-process flavour of :meth:`_alloc_vmm`: handles are used directly, no fds."""
        ll, ci = ctypes.c_longlong, ctypes.c_int
        world = len(comms)
        size, gran = comms[0]._vmm_size(nbytes)
        idx = [c.device.index or 0 for c in comms]
        mems = []
        for d in idx:
            h = ll()
            nv.call("vmm_create", d, size, 0, ctypes.byref(h))
            mems.append(h)
        mc = ll()
        nv.call("mc_create", world, size, 0, ctypes.byref(mc))
        for d in idx:
            nv.call("mc_add_device", mc, d)
        for h in mems:
            nv.call("mc_bind", mc, h, size)
        bufs = []
        for c, d in zip(comms, idx):
            # every GPU gets its own virtual ranges (W peer views + the multicast view), each
            # accessible from that GPU only, so a pointer always resolves to the GPU that uses it
            dev1 = (ci * 1)(d)
            ptrs = []
            for h in mems:
                out = ctypes.c_void_p()
                nv.call("vmm_map", h, size, gran, dev1, 1, ctypes.byref(out))
                ptrs.append(int(out.value))
            out = ctypes.c_void_p()
            nv.call("vmm_map", mc, size, gran, dev1, 1, ctypes.byref(out))
            # rank 0's buffer owns the handles
            own = (size, ptrs + [int(out.value)], (mems + [mc]) if c.rank == 0 else [])
            b = SymmBuffer(c, nbytes, ptrs[c.rank], ptrs, [], mc_ptr=int(out.value), vmm=own)
            with torch.cuda.device(c.device):
                b.bytes.zero_()
            c._buffers.append(b)
            bufs.append(b)
        for c in comms:
            torch.cuda.synchronize(c.device)
        return bufs

    @staticmethod
    def alloc_local(comms, nbytes: int):
        nbytes = (int(nbytes) + 255) // 256 * 256
        if comms[0].use_nvls and len(comms) > 1:
            return PeerComm._alloc_local_vmm(comms, nbytes)
        ptrs = []
        for c in comms:
            out = ctypes.c_void_p()
            with torch.cuda.device(c.device):
                nv.call("symm_alloc", nbytes, ctypes.byref(out))
            ptrs.append(int(out.value))
        bufs = []
        for c in comms:
            b = SymmBuffer(c, nbytes, ptrs[c.rank], ptrs, [])
            c._buffers.append(b)
            bufs.append(b)
        return bufs

    def multicast_supported(self) -> bool:
        """NVLS probe: can this GPU join a multicast object (``multimem`` collectives)?"""
        return nv.lib().dtdl_multicast_supported(self.device.index or 0) == 1

    # ------------------------------------------------------------------ collectives
    def _stream(self, stream, collective=True):
        """The stream a call runs on.  All collectives of one communicator share one pair of flag rows
        and device-resident epochs, so they must execute in issue order: when a collective is issued
        on another stream than the previous one, the new stream first waits for the old one.  (Not
        under stream capture, where a wait on outside work is illegal; the reducer keeps every
        collective of a captured step on its one communication stream.)"""
        s = stream or torch.cuda.current_stream(self.device)
        if collective:
            last = self._last_stream
            if last is not None and last != s and not torch.cuda.is_current_stream_capturing():
                s.wait_stream(last)
            self._last_stream = s
        return s.cuda_stream

    # Variant per message size, per world size, from the measured sweeps (profiles/round2/w{4,8}/
    # allreduce_sweep_*.md, profiles/allreduce_sweep_2gpu.md): rows of (largest message in bytes, variant),
    # first match wins; 0 = one-shot, 1 = two-shot, 2 = NVSwitch multicast (falls back to two-shot where
    # the buffers have no multicast mapping).  World sizes not listed use the next smaller entry.
    # DTDL_VARIANT_TABLE=/path/table.json ({"8": [[65536, 0], [null, 2]], ...}) replaces it.
    VARIANT_TABLE = {
        2: [(256 << 10, 0), (None, 1)],       # one-shot 20-23 us up to 256 KiB; two-shot wins above
        4: [(None, 2)],                        # multicast <= the others at every size measured
        8: [(None, 2)],
    }
    _variant_table = None

    @classmethod
    def variant_table(cls):
        if cls._variant_table is None:
            table = {w: list(rows) for w, rows in cls.VARIANT_TABLE.items()}
            path = os.environ.get("DTDL_VARIANT_TABLE")
            if path:
                import json
                with open(path) as f:
                    table = {int(w): [(None if lim is None else int(lim), int(v)) for lim, v in rows]
                             for w, rows in json.load(f).items()}
            cls._variant_table = table
        return cls._variant_table

    def choose_two_shot(self, nbytes: int) -> int:
        """The all-reduce variant for a message of ``nbytes``: 0 = one-shot, 1 = two-shot, 2 = multicast
        (looked up in ``VARIANT_TABLE``; multicast rows apply only when this communicator runs NVLS)."""
        if self.world <= 1:
            return 0
        table = self.variant_table()
        key = max((w for w in table if w <= self.world), default=min(table))
        for limit, variant in table[key]:
            if limit is None or nbytes <= limit:
                if variant == 2 and not (self.use_nvls and nbytes >= self.nvls_min_bytes):
                    variant = 1 if nbytes >= self.two_shot_min_bytes else 0
                return variant
        return 1

    def _nvls_ok(self, a: SymmBuffer, b: SymmBuffer, nbytes: int) -> bool:
        """Multicast path possible: both buffers are bound to a multicast object and the message
        is at least ``nvls_min_bytes``."""
        return bool(self.use_nvls and self.world > 1 and a.mc_ptr and b.mc_ptr
                    and nbytes >= self.nvls_min_bytes)

    def barrier(self, stream=None):
        nv.call("comm_barrier", self.handle, self._stream(stream))
        if self.world == 1:
            nv.add_launches(-1)

    def _count_sync(self):
        # flag-exchange kernels around the data kernel just launched: START always, DONE unless deferred
        if self.world > 1:
            nv.add_launches(1 if self._deferred else 2)

    def defer_done(self, on: bool):
        """``True``: the collectives that follow skip their trailing cross-GPU "done" exchange; close
        the sequence with ONE ``barrier()`` on the same stream before any rank touches the buffers
        again (what the gradient reducer does per training step).  ``False``: every collective is
        self-contained again."""
        self._deferred = bool(on)
        nv.lib().dtdl_comm_defer_done(self.handle, 1 if on else 0)

    # ------------------------------------------------------------------ device-side timeline
    TRACE_KINDS = {1: "wait(start)", 2: "wait(done)", 3: "reduce", 4: "nvls_reduce", 5: "broadcast",
                   6: "allgather", 7: "mark"}

    def trace_begin(self, capacity: int = 256):
        """Every sync / data kernel launched through this communicator from now on stamps its first
        block's start and last block's end (``%globaltimer``) into a device buffer -- also when the
        launches are replayed from a CUDA graph.  Call before the step is captured / run."""
        with torch.cuda.device(self.device):
            self._trace = torch.zeros((capacity, 4), dtype=torch.int64, device=self.device)
        nv.lib().dtdl_comm_trace(self.handle, ctypes.c_void_p(self._trace.data_ptr()), capacity)

    def trace_mark(self, stream=None):
        """A time stamp on ``stream`` (e.g. right after the last backward kernel)."""
        nv.call("comm_trace_mark", self.handle, 7, self._stream(stream, collective=False))

    def trace_end(self):
        nv.lib().dtdl_comm_trace(self.handle, None, 0)

    def trace_rows(self):
        """[(kind, t0_us, t1_us, bytes)] of the last execution of the traced launches, times relative
        to the earliest start."""
        if getattr(self, "_trace", None) is None:
            return []
        torch.cuda.synchronize(self.device)
        t = self._trace.cpu()
        rows = [(int(k), int(a), int(b), int(n)) for a, b, k, n in t.tolist() if k]
        if not rows:
            return []
        t_min = min(a for _, a, _, _ in rows)
        return [(self.TRACE_KINDS.get(k, str(k)), (a - t_min) / 1e3, (b - t_min) / 1e3, n) for k, a, b, n in rows]

    def allreduce(self, buf: SymmBuffer, offset_bytes: int, numel: int, scale: float = 1.0,
                  two_shot=None, dst: SymmBuffer = None, stream=None, max_blocks=None):
        """In-place (or into ``dst``) sum * scale of ``numel`` fp32 elements on every rank."""
        in_place = dst is None or dst is buf
        dst = dst or buf
        ts = self.choose_two_shot(numel * 4) if two_shot is None else int(two_shot)
        if ts == 2 and not self._nvls_ok(buf, dst, numel * 4):
            ts = 1
        if in_place and self.world > 1 and ts == 0:
            # one-shot reads every rank's whole buffer, so it may not overwrite its source;
            # two-shot and multicast only write the slice that nobody else reads
            ts = 1
        if ts == 2:
            nv.call("comm_nvls_allreduce", self.handle, ctypes.c_void_p(buf.mc_ptr), ctypes.c_void_p(dst.mc_ptr),
                    offset_bytes, numel, scale, max_blocks or self.max_blocks, self._stream(stream))
            self._count_sync()
            return
        nv.call("comm_allreduce", self.handle, buf.ptr_array, dst.ptr_array, offset_bytes, numel, scale, ts,
                max_blocks or self.max_blocks, self._stream(stream))
        self._count_sync()

    def reduce_update(self, grads: SymmBuffer, params: SymmBuffer, mom, var, hyper, offset_bytes: int,
                      numel: int, opt: int, two_shot=None, stream=None, max_blocks=None):
        """Fused gradient all-reduce + 1/W + weight decay + SGD/Adam update of a bucket."""
        ts = self.choose_two_shot(numel * 4) if two_shot is None else int(two_shot)
        if ts == 2 and not self._nvls_ok(grads, params, numel * 4):
            ts = 1
        if ts == 2:
            nv.call("comm_nvls_reduce_update", self.handle, ctypes.c_void_p(grads.mc_ptr),
                    ctypes.c_void_p(params.mc_ptr), ctypes.c_void_p(params.local_ptr), nv.ptr(mom), nv.ptr(var),
                    nv.ptr(hyper), offset_bytes, numel, opt, max_blocks or self.max_blocks, self._stream(stream))
            self._count_sync()
            return
        nv.call("comm_reduce_update", self.handle, grads.ptr_array, params.ptr_array, nv.ptr(mom),
                nv.ptr(var), nv.ptr(hyper), offset_bytes, numel, opt, ts, max_blocks or self.max_blocks,
                self._stream(stream))
        self._count_sync()

    def reduce_update16(self, grads16: SymmBuffer, params: SymmBuffer, mom, var, hyper, offset_bytes: int,
                        numel: int, opt: int, stream=None, max_blocks=None):
        """``reduce_update`` with the gradients travelling as bf16 (``grads16`` = the bf16 gradient arena,
        already filled by ``cast_f32_bf16``): half the NVLink bytes, fp32 accumulation and update.
        Multicast variant when both buffers are bound to a multicast object, else two-shot."""
        if self._nvls_ok(grads16, params, numel * 2):
            nv.call("comm_nvls_reduce_update16", self.handle, ctypes.c_void_p(grads16.mc_ptr),
                    ctypes.c_void_p(params.mc_ptr), ctypes.c_void_p(params.local_ptr), nv.ptr(mom), nv.ptr(var),
                    nv.ptr(hyper), offset_bytes, numel, opt, max_blocks or self.max_blocks, self._stream(stream))
        else:
            nv.call("comm_reduce_update16", self.handle, grads16.ptr_array, params.ptr_array, nv.ptr(mom),
                    nv.ptr(var), nv.ptr(hyper), offset_bytes, numel, opt, max_blocks or self.max_blocks,
                    self._stream(stream))
        self._count_sync()

    def broadcast(self, buf: SymmBuffer, offset_bytes: int, nbytes: int, root: int = 0, stream=None,
                  max_blocks=None):
        if self._nvls_ok(buf, buf, nbytes):
            nv.call("comm_nvls_broadcast", self.handle, ctypes.c_void_p(buf.local_ptr), ctypes.c_void_p(buf.mc_ptr),
                    offset_bytes, nbytes, root, max_blocks or self.max_blocks, self._stream(stream))
            self._count_sync()
            return
        nv.call("comm_broadcast", self.handle, buf.ptr_array, offset_bytes, nbytes, root,
                max_blocks or self.max_blocks, self._stream(stream))
        self._count_sync()

    def allgather(self, buf: SymmBuffer, offset_bytes: int, nbytes: int, stream=None, max_blocks=None):
        nv.call("comm_allgather", self.handle, buf.ptr_array, offset_bytes, nbytes,
                max_blocks or self.max_blocks, self._stream(stream))
        self._count_sync()

    def close(self):
        torch.cuda.synchronize(self.device)
        for b in self._buffers:
            b.free()
        self._buffers = []
        if self._fdchan is not None:
            self._fdchan.close()
            self._fdchan = None
        if self.handle:
            nv.lib().dtdl_comm_destroy(self.handle)
            self.handle = ctypes.c_void_p()
