from dgraph_b200.utils.data_splitting import largest_split, split_per_rank, try_barrier
from dgraph_b200.utils.timing import TimingReport

__all__ = ["largest_split", "split_per_rank", "try_barrier", "TimingReport"]
