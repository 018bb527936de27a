"""Distributed helpers of the GraphCast experiment (parity: reference ``experiments/GraphCast/dist_utils.py:8-129``).

* :func:`SingleProcessDummyCommunicator` -- a working single-process communicator (the reference's class of that name cannot
  be instantiated); here it is a :class:`Communicator` bound to the world-size-1 engine.
* :class:`CommAwareDistributedSampler` -- every rank of one graph replica sees the same sample index, different replicas see
  different samples (the reference's ``__iter__`` returns ``range(len(dataset))`` for everyone).
"""
from __future__ import annotations

import math
from typing import Iterator

import torch
from torch.utils.data import Sampler


def _communicator_class():
    from dgraph_b200.communicator import Communicator

    return Communicator


class SingleProcessDummyCommunicator(_communicator_class()):
    """World-size-1 communicator for debugging and single-GPU runs: a full :class:`Communicator` on the local engine (every
    gather / scatter / halo call works and costs no communication) that does not claim the process-wide singleton."""

    def __init__(self):
        from dgraph_b200.parallel.local_engine import LocalEngine

        super().__init__("single", _engine=LocalEngine())
        self._ranks_per_sample = 1

    def __str__(self) -> str:
        return self.backend

    def rank_cuda_device(self):
        return torch.cuda.current_device()


class CommAwareDistributedSampler(Sampler):
    def __init__(self, dataset, comm=None, shuffle: bool = True, seed: int = 0, drop_last: bool = False,
                 num_replicas: int = None, replica_id: int = None):
        self.dataset = dataset
        self.num_replicas = num_replicas if num_replicas is not None else (comm.get_num_replicas() if comm is not None else 1)
        self.replica_id = replica_id if replica_id is not None else (comm.get_replica_id() if comm is not None else 0)
        self.shuffle, self.seed, self.epoch, self.drop_last = shuffle, seed, 0, drop_last
        n = len(dataset)
        self.num_samples = n // self.num_replicas if drop_last else math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def set_epoch(self, epoch: int):
        self.epoch = epoch

    def __len__(self) -> int:
        return self.num_samples

    def __iter__(self) -> Iterator[int]:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            idx = torch.randperm(n, generator=g).tolist()
        else:
            idx = list(range(n))
        if not self.drop_last and len(idx) < self.total_size:
            idx += idx[: self.total_size - len(idx)]
        idx = idx[: self.total_size]
        return iter(idx[self.replica_id:self.total_size:self.num_replicas])
