"""b200stage: B200-native W8A8 decoder-block stage path.

Importable as ``importlib.import_module("llm-for-distributed-egde-devices_b200")`` or through the
``b200stage`` alias module at the repository root.
"""
from .config import ModelConfig, PRESETS, get_config  # noqa: F401
from . import _abi  # noqa: F401

__all__ = ["ModelConfig", "PRESETS", "get_config", "Stage", "KVCache", "Pipeline", "DistPipeline", "Ensemble", "build"]


def build(verbose: bool = False):
    return _abi.build(verbose)


def __getattr__(name):   # lazy: stage/pipeline import torch
    if name in ("Stage", "KVCache"):
        from . import stage
        return getattr(stage, name)
    if name in ("Pipeline", "DistPipeline"):
        from . import pipeline
        return getattr(pipeline, name)
    if name == "Ensemble":
        from .ensemble import Ensemble
        return Ensemble
    if name in ("weights", "ops", "stage", "pipeline", "rpc", "config", "loader", "ensemble"):     # b200stage.<submodule> without a prior import
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
