"""Process-group bootstrap: ``tcp://``, ``env://``, SLURM and host-list rendezvous.

The reference hand-types ``--init-method tcp://host:port --rank r --world-size w`` per shell
(pytorch/distributed_data_parallel.py:45-48,61-62), builds a ``TF_CONFIG`` from
``--worker_hosts`` / ``--task_index`` (tensorflow2/mnist_multi_worker_strategy.py:18-24) and
gets ranks from MPI under ChainerMN (chainer/train_mnist_multi.py:55-56).  ``init`` accepts all
of those spellings plus ``env://`` (torchrun) and SLURM's ``srun`` environment, and returns a
:class:`Communicator` exposing ``rank`` / ``size`` / ``intra_rank`` like ChainerMN's handle.

Backends
    ``"b200"``  GPU data plane = this package's peer-memory kernels over NVLink (no NCCL);
                control plane = a gloo group.  Default when CUDA is available.
    ``"nccl"``  stock NCCL collectives (the baseline configuration, kept for comparison).
    ``"gloo"``  CPU tensors (plumbing tests, config 1 of BASELINE.json).
"""
from __future__ import annotations

import datetime
import os
import re
from dataclasses import dataclass, field

import torch
import torch.distributed as dist

_DEFAULT_PORT = 29500


# --------------------------------------------------------------------------------------
# spec parsing (pure functions: unit-tested without any process group)
# --------------------------------------------------------------------------------------
@dataclass
class RendezvousSpec:
    master_addr: str = "127.0.0.1"
    master_port: int = _DEFAULT_PORT
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    source: str = "single"

    @property
    def init_method(self) -> str:
        return f"tcp://{self.master_addr}:{self.master_port}"


def expand_slurm_nodelist(nodelist: str):
    """``"gpu[01-03,07],login1"`` -> ``["gpu01","gpu02","gpu03","gpu07","login1"]``."""
    hosts, i, n = [], 0, len(nodelist)
    while i < n:
        j, depth = i, 0
        while j < n and (nodelist[j] != "," or depth):
            depth += nodelist[j] == "["
            depth -= nodelist[j] == "]"
            j += 1
        token = nodelist[i:j]
        i = j + 1
        m = re.fullmatch(r"([^\[]*)\[([^\]]*)\](.*)", token)
        if not m:
            if token:
                hosts.append(token)
            continue
        prefix, body, suffix = m.groups()
        for part in body.split(","):
            if "-" in part:
                lo, hi = part.split("-")
                width = len(lo)
                for v in range(int(lo), int(hi) + 1):
                    hosts.append(f"{prefix}{v:0{width}d}{suffix}")
            else:
                hosts.append(f"{prefix}{part}{suffix}")
    return hosts


def spec_from_env(env=None) -> RendezvousSpec:
    """torchrun / ``env://`` variables (RANK, WORLD_SIZE, MASTER_ADDR, MASTER_PORT, LOCAL_RANK)."""
    env = os.environ if env is None else env
    missing = [k for k in ("RANK", "WORLD_SIZE") if k not in env]
    if missing:
        raise ValueError(f"env:// rendezvous needs {missing} in the environment")
    return RendezvousSpec(
        master_addr=env.get("MASTER_ADDR", "127.0.0.1"),
        master_port=int(env.get("MASTER_PORT", _DEFAULT_PORT)),
        rank=int(env["RANK"]), world_size=int(env["WORLD_SIZE"]),
        local_rank=int(env.get("LOCAL_RANK", env["RANK"])), source="env")


def spec_from_slurm(env=None, port=None) -> RendezvousSpec:
    """``srun`` variables: SLURM_PROCID / SLURM_NTASKS / SLURM_LOCALID / SLURM_NODELIST."""
    env = os.environ if env is None else env
    if "SLURM_PROCID" not in env or "SLURM_NTASKS" not in env:
        raise ValueError("SLURM rendezvous needs SLURM_PROCID and SLURM_NTASKS")
    nodelist = env.get("SLURM_STEP_NODELIST") or env.get("SLURM_JOB_NODELIST") or env.get("SLURM_NODELIST")
    master = env.get("MASTER_ADDR") or env.get("SLURM_LAUNCH_NODE_IPADDR")
    if nodelist:
        master = env.get("MASTER_ADDR") or expand_slurm_nodelist(nodelist)[0]
    if port is None:
        # derive a port from the job id so concurrent jobs on one node do not collide
        port = int(env.get("MASTER_PORT", 15000 + int(env.get("SLURM_JOB_ID", "0")) % 20000))
    return RendezvousSpec(master_addr=master or "127.0.0.1", master_port=int(port),
                          rank=int(env["SLURM_PROCID"]), world_size=int(env["SLURM_NTASKS"]),
                          local_rank=int(env.get("SLURM_LOCALID", 0)), source="slurm")


def spec_from_hosts(worker_hosts, task_index: int) -> RendezvousSpec:
    """``--worker_hosts a:p,b:p --task_index i`` (the reference's TF_CONFIG spelling)."""
    hosts = worker_hosts.split(",") if isinstance(worker_hosts, str) else list(worker_hosts)
    hosts = [h.strip() for h in hosts if h.strip()]
    if not 0 <= task_index < len(hosts):
        raise ValueError(f"task_index {task_index} outside worker list of {len(hosts)}")
    addr, _, port = hosts[0].partition(":")
    mine = hosts[task_index].partition(":")[0]
    local_rank = sum(1 for h in hosts[:task_index] if h.partition(":")[0] == mine)
    return RendezvousSpec(master_addr=addr, master_port=int(port or _DEFAULT_PORT), rank=task_index,
                          world_size=len(hosts), local_rank=local_rank, source="hosts")


def resolve_spec(init_method=None, rank=None, world_size=None, local_rank=None, worker_hosts=None,
                 task_index=None, env=None) -> RendezvousSpec:
    env = os.environ if env is None else env
    if worker_hosts is not None:
        spec = spec_from_hosts(worker_hosts, int(task_index or 0))
    elif init_method is None:
        if "RANK" in env and "WORLD_SIZE" in env:
            spec = spec_from_env(env)
        elif "SLURM_PROCID" in env and "SLURM_NTASKS" in env:
            spec = spec_from_slurm(env)
        else:
            spec = RendezvousSpec(rank=rank or 0, world_size=world_size or 1)
    elif init_method.startswith("env://"):
        spec = spec_from_env(env)
    elif init_method.startswith("slurm"):
        spec = spec_from_slurm(env)
    elif init_method.startswith("tcp://"):
        hostport = init_method[len("tcp://"):]
        addr, _, port = hostport.rpartition(":")
        if rank is None or world_size is None:
            raise ValueError("tcp:// rendezvous needs explicit rank and world_size "
                             "(pytorch/distributed_data_parallel.py:47-48)")
        spec = RendezvousSpec(master_addr=addr, master_port=int(port), rank=int(rank),
                              world_size=int(world_size), source="tcp")
    else:
        raise ValueError(f"unsupported init_method {init_method!r}")
    if rank is not None and spec.source in ("single", "tcp"):
        spec.rank = int(rank)
    if world_size is not None and spec.source in ("single", "tcp"):
        spec.world_size = int(world_size)
    if local_rank is not None:
        spec.local_rank = int(local_rank)
    elif spec.source == "tcp":
        spec.local_rank = int(env.get("LOCAL_RANK", 0))
    return spec


# --------------------------------------------------------------------------------------
# communicator
# --------------------------------------------------------------------------------------
@dataclass
class Communicator:
    """Rank handle (ChainerMN's ``comm.rank / size / intra_rank``, chainer/train_mnist_multi.py:55-66)."""
    rank: int
    size: int
    intra_rank: int
    backend: str
    device: torch.device
    group: object = None          # torch.distributed group carrying tensors of `device`
    control: object = None        # CPU (gloo) group used for objects / bootstrap
    peer: object = None           # PeerComm for the "b200" backend
    spec: RendezvousSpec = field(default_factory=RendezvousSpec)

    # aliases
    @property
    def world_size(self):
        return self.size

    @property
    def local_rank(self):
        return self.intra_rank

    def barrier(self):
        if self.size > 1:
            dist.barrier(group=self.control)

    def broadcast_object(self, obj, src=0):
        if self.size == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src, group=self.control)
        return box[0]

    def allgather_object(self, obj):
        if self.size == 1:
            return [obj]
        out = [None] * self.size
        dist.all_gather_object(out, obj, group=self.control)
        return out

    def allreduce_mean(self, value: float) -> float:
        """Mean of a python scalar over ranks (multi-node evaluator, :102-104)."""
        if self.size == 1:
            return float(value)
        t = torch.tensor([float(value)], dtype=torch.float64)
        dist.all_reduce(t, group=self.control)
        return float(t.item() / self.size)


_current: Communicator | None = None


def current() -> Communicator | None:
    return _current


def init(backend=None, init_method=None, rank=None, world_size=None, local_rank=None,
         worker_hosts=None, task_index=None, device=None, timeout_s=600) -> Communicator:
    """Joins the job and returns the :class:`Communicator`.  Idempotent per process."""
    global _current
    if _current is not None:
        return _current
    spec = resolve_spec(init_method, rank, world_size, local_rank, worker_hosts, task_index)
    cuda = torch.cuda.is_available()
    if backend is None:
        backend = "b200" if cuda else "gloo"
    backend = {"peer": "b200", "pure_nccl": "b200", "naive": "gloo"}.get(backend, backend)
    if backend in ("b200", "nccl") and not cuda:
        raise RuntimeError(f"backend {backend!r} needs a CUDA device; use backend='gloo' on CPU")
    if device is None:
        if backend == "gloo" and not cuda:
            device = torch.device("cpu")
        elif backend == "gloo":
            device = torch.device("cpu")
        else:
            ndev = torch.cuda.device_count()
            device = torch.device("cuda", spec.local_rank % max(ndev, 1))
    device = torch.device(device)
    if device.type == "cuda":
        torch.cuda.set_device(device)

    group = control = None
    if spec.world_size > 1 or dist.is_initialized():
        if not dist.is_initialized():
            pg_backend = "nccl" if backend == "nccl" else "gloo"
            dist.init_process_group(backend=pg_backend, init_method=spec.init_method, rank=spec.rank,
                                    world_size=spec.world_size,
                                    timeout=datetime.timedelta(seconds=timeout_s))
        group = dist.group.WORLD
        control = group if dist.get_backend() == "gloo" else dist.new_group(backend="gloo")
    comm = Communicator(rank=spec.rank, size=spec.world_size, intra_rank=spec.local_rank, backend=backend,
                        device=device, group=group, control=control, spec=spec)
    if backend == "b200":
        from .peer import PeerComm
        comm.peer = PeerComm.create_multiprocess(comm)
    _current = comm
    return comm


def shutdown():
    global _current
    if _current is not None and _current.peer is not None:
        if _current.size > 1 and dist.is_initialized():
            # nobody unmaps memory a slower rank might still be touching
            try:
                torch.cuda.synchronize(_current.device)
                dist.monitored_barrier(group=_current.control, timeout=datetime.timedelta(seconds=60))
            except Exception:   # noqa: BLE001 - a peer that already died must not block our exit
                pass
        _current.peer.close()
    _current = None
    if dist.is_initialized():
        dist.destroy_process_group()


# ChainerMN spelling (chainer/train_mnist_multi.py:55,61)
def create_communicator(name="pure_nccl", **kw) -> Communicator:
    return init(backend=name, **kw)
