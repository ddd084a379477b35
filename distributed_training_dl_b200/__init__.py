"""B200-native data-parallel training framework (see README.md / DESIGN.md).

    import distributed_training_dl_b200 as dtdl
    comm = dtdl.init()                       # tcp:// | env:// | SLURM | host list
    net = dtdl.wrap(dtdl.models.pyramidnet().cuda())
    opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    loss = criterion(net(x), y); opt.zero_grad(); loss.backward(); opt.step()
"""
__version__ = "0.1.0"

from . import _native  # noqa: F401
from . import models, ops, parallel, utils  # noqa: F401
from .engine import CrossEntropyLoss, TrainStep
from .ops import set_precision
from .parallel import (DataParallel, DataParallelStep, FusedAdam, FusedSGD, DistributedDataParallel, DistributedModel,
                       create_communicator, current, init, shutdown, wrap)
from .utils import (DistributedSampler, SyntheticDataset, load_model, load_snapshot, save_model,
                    save_snapshot, scatter_dataset, spawn)
