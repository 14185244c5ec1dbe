"""fit-slam_b200: B200-native FIT-SLAM exploration hot path (C-ABI library + host mirrors).

The directory name has a hyphen, so load it with `importlib` (see tests/conftest.py) under the
module name `fitslam_b200`.
"""
from . import capi  # noqa: F401
from .capi import Context, FitslamError, Params, default_params, load_library  # noqa: F401
