"""Reference file name (``experiments/GraphCast/graphcast_config.py:18-88``)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.models.graphcast import Config, DataConfig, ModelConfig, TrainingConfig  # noqa: F401,E402

__all__ = ["TrainingConfig", "DataConfig", "ModelConfig", "Config"]
