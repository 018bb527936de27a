"""Abstract communicator contract.

Parity: reference ``DGraph/CommunicatorBase.py:17-50`` (same abstract method set and the
class-level ``_is_initialized`` flag).
"""
from abc import ABC, abstractmethod


class CommunicatorBase(ABC):
    _is_initialized = False

    def __init__(self):
        self.backend = ""

    @abstractmethod
    def init_process_group(self, backend: str, **kwargs):
        ...

    @abstractmethod
    def get_rank(self) -> int:
        ...

    @abstractmethod
    def get_world_size(self) -> int:
        ...

    @abstractmethod
    def barrier(self):
        ...

    @abstractmethod
    def scatter(self, *args, **kwargs):
        ...

    @abstractmethod
    def gather(self, *args, **kwargs):
        ...

    @abstractmethod
    def destroy(self):
        ...
