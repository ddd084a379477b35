from .arena import BufferArena, ParamArena
from .optim import FusedAdam, FusedSGD
from .rendezvous import (Communicator, RendezvousSpec, create_communicator, current,
                         expand_slurm_nodelist, init, resolve_spec, shutdown, spec_from_env,
                         spec_from_hosts, spec_from_slurm)
from .wrap import DistributedDataParallel, DistributedModel, Reducer, wrap
from .data_parallel import DataParallel, DataParallelStep
