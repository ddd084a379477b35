from .checkpoint import (add_module_prefix, load_model, load_snapshot, save_model, save_snapshot,
                         strip_module_prefix)
from .data import DistributedSampler, PinnedPrefetcher, SyntheticDataset, scatter_dataset
from .launch import free_port, spawn
from .logging import (JsonLog, RankLogger, TrainMeter, format_bucket_report, format_elapsed, get_rank,
                      rank0_only)
from .timing import ClockSampler, DeviceTimer, flush_l2, max_over_ranks
