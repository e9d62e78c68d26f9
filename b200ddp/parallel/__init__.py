from .buckets import build_layout, plan_buckets  # noqa: F401
from .process_group import cleanup_distributed, get_device, setup_distributed  # noqa: F401


def __getattr__(name):
    if name == "DataParallel":
        from .engine import DataParallel
        return DataParallel
    raise AttributeError(name)
