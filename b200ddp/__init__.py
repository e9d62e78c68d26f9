"""b200ddp — a Blackwell-native data-parallel trainer.

Same capabilities as northflank-examples/distributed-pytorch-example (launcher,
env:// rendezvous, sharded sampling, train loop with rank-0 logging, compatible
checkpoints); the hot path is hand-written sm_100a CUDA.  See DESIGN.md.
"""
from .checkpoint import load_checkpoint, save_checkpoint  # noqa: F401
from .data import BatchLoader, ShardSampler, SyntheticDataset, create_data_loader  # noqa: F401
from .models.simplenet import SimpleNet  # noqa: F401
from .optim import FlatOptimizer  # noqa: F401
from .parallel.engine import DataParallel  # noqa: F401
from .parallel.process_group import cleanup_distributed, get_device, setup_distributed  # noqa: F401

__version__ = "0.1.0"
