"""Configs of the RGAT experiment (parity: reference ``experiments/OGB-LSC/config.py:18-45``)."""
from dataclasses import dataclass


@dataclass
class ModelConfig:
    hidden_channels: int = 64  # reference default is 2 (a debugging value); 64 keeps rows 16-byte multiples for the kernels
    dropout: float = 0.5
    num_layers: int = 2
    heads: int = 4
    use_cache: bool = True


@dataclass
class TrainingConfig:
    epochs: int = 100
    lr: float = 1e-4
    weight_decay: float = 5e-4
    lr_step_size: int = 25
    lr_gamma: float = 0.25


@dataclass
class SyntheticDatasetConfig:
    num_papers: int = 32768
    num_authors: int = 16384
    num_institutions: int = 256
    num_features: int = 16
    num_classes: int = 153
