"""B200-native batched implementation of the ExoskeletonEnv_train reset()/step() hot path.

The importable alias of this package is `exo_b200` (the directory name carries hyphens)."""
from ._capi import ExoError, load as load_library  # noqa: F401
from .lap import LAP  # noqa: F401
from .policy import DevicePolicy, load_reference_policy  # noqa: F401
from .td7 import DeviceTD7, Hyperparameters  # noqa: F401
from .train import train_device, pink_noise  # noqa: F401
from .evaluate import evaluate_device  # noqa: F401
from .env import BatchedExoskeletonEnv, ExoskeletonEnv_train, INFO_KEYS, REWARD_KEYS, STAT_NAMES  # noqa: F401

__all__ = ["LAP", "DevicePolicy", "load_reference_policy", "DeviceTD7", "Hyperparameters", "train_device", "pink_noise", "evaluate_device", "BatchedExoskeletonEnv", "ExoskeletonEnv_train", "ExoError", "load_library", "INFO_KEYS", "REWARD_KEYS", "STAT_NAMES"]
