"""Alias package: ``k8s_distributed_deeplearning_b200`` is ``b200dp`` under the
upstream repository's name (``models/ ops/ parallel/ utils/`` live in ``b200dp/``)."""
import importlib
import sys

import b200dp as _impl
from b200dp import *  # noqa: F401,F403

for _sub in ("models", "ops", "parallel", "utils", "runtime", "train", "ckpt", "launch", "data", "callbacks"):
    _m = importlib.import_module(f"b200dp.{_sub}")
    sys.modules[f"{__name__}.{_sub}"] = _m
    globals()[_sub] = _m
__version__ = _impl.__version__
