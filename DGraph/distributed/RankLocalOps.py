from dgraph_b200.parallel.rank_local import *  # noqa: F401,F403
from dgraph_b200.parallel.rank_local import (  # noqa: F401
    OptimizedLocalScatterGather, OptimizedLocalScatterSumGather, OptimizedRankLocalMaskedGather, RankLocalMaskedGather,
    RankLocalMaskedScatter, RankLocalRenumbering, RankLocalRenumberingWithMapping, LocalAggregateWithRemapping,
)
