from .mcts_ctree import (EfficientZeroMCTSCtree, GumbelMuZeroMCTSCtree, MuZeroMCTSCtree,  # noqa: F401
                         MuZeroRNNFullObsMCTSCtree, UniZeroMCTSCtree)
from .mcts_ctree_sampled import SampledEfficientZeroMCTSCtree  # noqa: F401
from .mcts_ctree_stochastic import StochasticMuZeroMCTSCtree  # noqa: F401
