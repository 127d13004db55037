"""B200-native batched MCTS for LightZero: the reference's tree API over hand-written sm_100a CUDA.

    from lightzero_b200 import MuZeroMCTSCtree as MCTSCtree      # instead of `from lzero.mcts import ...`

Importing the package does not need a GPU; creating roots or searching does (there is no CPU fallback)."""
from .tree_search import (EfficientZeroMCTSCtree, GumbelMuZeroMCTSCtree, MuZeroMCTSCtree,  # noqa: F401
                          MuZeroRNNFullObsMCTSCtree, SampledEfficientZeroMCTSCtree, StochasticMuZeroMCTSCtree,
                          UniZeroMCTSCtree)
from . import callers  # noqa: F401,E402  (collect_search / reanalyze_policy_targets / reanalyze_value_targets)
