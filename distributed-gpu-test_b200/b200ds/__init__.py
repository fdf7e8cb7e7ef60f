"""b200ds -- the names train.py imports from `deepspeed` (train.py:11-12,59,61,87), backed
by the B200-native C-ABI library libb200ds.so.  Use as `import b200ds as deepspeed`.

Scope: the one data-parallel hot path of dashabalashova/distributed-gpu-test (the
synthetic-data training step of train.py); nothing else of DeepSpeed is provided.
"""
from . import _lib, accelerator, comm  # noqa: F401
from .accelerator import get_accelerator  # noqa: F401
from .comm import init_distributed  # noqa: F401
from .data import DeviceBatchLoader  # noqa: F401
from .engine import B200Engine, BCEWithLogitsLoss  # noqa: F401

__version__ = "0.1.0"


def initialize(args=None, model=None, optimizer=None, model_parameters=None, training_data=None,
               lr_scheduler=None, config=None, config_params=None, **_ignored):
    """Drop-in for deepspeed.initialize as called at train.py:87-93.  `args` may be the
    config dict itself (the reference passes ds_config as both args= and config=)."""
    if model is None:
        raise ValueError("initialize: model is required")
    if config is None:
        config = config_params if config_params is not None else (args if isinstance(args, dict) else None)
    if optimizer is None:
        raise NotImplementedError("initialize: pass the client optimizer, as train.py:85,90 does")
    if lr_scheduler is not None or training_data is not None:
        raise NotImplementedError("initialize: lr_scheduler/training_data are not part of the reference path")
    engine = B200Engine(model, optimizer, config)
    return engine, optimizer, None, None
