"""Reference file name (``lsc_datasets/synthetic_dataset.py:29-199``)."""
from . import SyntheticHeterogeneousDataset  # noqa: F401
