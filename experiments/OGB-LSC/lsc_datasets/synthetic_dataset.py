"""Reference file name (``lsc_datasets/synthetic_dataset.py:29-199``)."""
from . import SyntheticHeterogeneousDataset  # noqa: F401

import hashlib  # noqa: E402

import torch  # noqa: E402


def generate_config_hash(numbers) -> str:
    """Stable name for a dataset configuration (plan-cache file names): md5 of the tuple's text form."""
    return hashlib.md5(str(tuple(numbers)).encode("utf-8")).hexdigest()


def _random_relation(num_src: int, num_dst: int, num_edges: int, generator=None) -> torch.Tensor:
    """``[2, E']`` distinct random (source, destination) pairs, at most ``num_edges`` of them."""
    num_edges = max(int(num_edges), 1)
    e = torch.stack([torch.randint(0, num_src, (num_edges,), generator=generator),
                     torch.randint(0, num_dst, (num_edges,), generator=generator)])
    return torch.unique(e, dim=1)


def _generate_paper_2_paper_edges(num_papers: int, generator=None) -> torch.Tensor:
    """Citations, about 11 per paper (the MAG240M average)."""
    return _random_relation(num_papers, num_papers, num_papers * 11, generator)


def _generate_author_2_paper_edges(num_authors: int, num_papers: int, generator=None) -> torch.Tensor:
    """Authorship, about 3.5 papers per author."""
    return _random_relation(num_authors, num_papers, int(num_authors * 3.5), generator)


def _generate_author_2_institution_edges(num_authors: int, num_institutions: int, generator=None) -> torch.Tensor:
    """Affiliations, about 0.35 per author."""
    return _random_relation(num_authors, num_institutions, int(num_authors * 0.35), generator)
