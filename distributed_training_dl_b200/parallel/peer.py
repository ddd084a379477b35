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
"""
from __future__ import annotations

import ctypes

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
    t._dtdl_raw = raw  # keep the holder (and thereby the allocation owner) alive
    return t


class SymmBuffer:
    """Same-sized allocation on every rank + the per-rank base pointer table."""

    def __init__(self, comm: "PeerComm", nbytes: int, local_ptr: int, peer_ptrs, opened):
        self.comm, self.nbytes, self.local_ptr = comm, nbytes, local_ptr
        self.peer_ptrs = list(peer_ptrs)
        self._opened = opened  # IPC mappings to close
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
        self.max_blocks = 32        # SM budget of a collective kernel (leaves SMs to backward)
        self.two_shot_min_bytes = 256 * 1024

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
        return comms

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

    def alloc(self, nbytes: int, tag: str = "buf") -> SymmBuffer:
        """Collective: every rank must call with the same size in the same order."""
        nbytes = (int(nbytes) + 255) // 256 * 256
        if self._local_group is not None:
            raise RuntimeError("use PeerComm.alloc_local(comms, nbytes) in single-process mode")
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
    def alloc_local(comms, nbytes: int):
        nbytes = (int(nbytes) + 255) // 256 * 256
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
    def _stream(self, stream):
        return (stream or torch.cuda.current_stream(self.device)).cuda_stream

    def choose_two_shot(self, nbytes: int) -> int:
        """One-shot below the measured crossover, two-shot above (see profiles/allreduce_sweep.md)."""
        return 1 if (self.world > 1 and nbytes >= self.two_shot_min_bytes) else 0

    def barrier(self, stream=None):
        nv.call("comm_barrier", self.handle, self._stream(stream))

    def allreduce(self, buf: SymmBuffer, offset_bytes: int, numel: int, scale: float = 1.0,
                  two_shot=None, dst: SymmBuffer = None, stream=None, max_blocks=None):
        """In-place (or into ``dst``) sum * scale of ``numel`` fp32 elements on every rank."""
        in_place = dst is None or dst is buf
        dst = dst or buf
        ts = self.choose_two_shot(numel * 4) if two_shot is None else int(two_shot)
        if in_place and self.world > 1:
            # one-shot reads every rank's whole buffer, so it may not overwrite its source;
            # two-shot only writes the slice that nobody else reads
            ts = 1
        nv.call("comm_allreduce", self.handle, buf.ptr_array, dst.ptr_array, offset_bytes, numel, scale, ts,
                max_blocks or self.max_blocks, self._stream(stream))

    def reduce_update(self, grads: SymmBuffer, params: SymmBuffer, mom, var, hyper, offset_bytes: int,
                      numel: int, opt: int, two_shot=None, stream=None, max_blocks=None):
        """Fused gradient all-reduce + 1/W + weight decay + SGD/Adam update of a bucket."""
        ts = self.choose_two_shot(numel * 4) if two_shot is None else int(two_shot)
        nv.call("comm_reduce_update", self.handle, grads.ptr_array, params.ptr_array, nv.ptr(mom),
                nv.ptr(var), nv.ptr(hyper), offset_bytes, numel, opt, ts, max_blocks or self.max_blocks,
                self._stream(stream))

    def broadcast(self, buf: SymmBuffer, offset_bytes: int, nbytes: int, root: int = 0, stream=None,
                  max_blocks=None):
        nv.call("comm_broadcast", self.handle, buf.ptr_array, offset_bytes, nbytes, root,
                max_blocks or self.max_blocks, self._stream(stream))

    def allgather(self, buf: SymmBuffer, offset_bytes: int, nbytes: int, stream=None, max_blocks=None):
        nv.call("comm_allgather", self.handle, buf.ptr_array, offset_bytes, nbytes,
                max_blocks or self.max_blocks, self._stream(stream))

    def close(self):
        torch.cuda.synchronize(self.device)
        for b in self._buffers:
            b.free()
        self._buffers = []
        if self.handle:
            nv.lib().dtdl_comm_destroy(self.handle)
            self.handle = ctypes.c_void_p()
