"""Drop-in for lzero/mcts/ctree/ctree_muzero/mz_tree.pyx over the CUDA forest (variant LZB_MZ)."""
from .._tree_api import (MinMaxStatsList, NodeBase, ResultsWrapper, RootsBase, batch_backpropagate,  # noqa: F401
                         batch_backpropagate_with_reuse, batch_traverse, batch_traverse_with_reuse, set_tie_break)


class Roots(RootsBase):
    """mz_tree.pyx:26-58: Roots(root_num, legal_actions_list)."""
    variant = "mz"


class Node(NodeBase):
    """mz_tree.pyx:60-72."""
    _roots_cls = Roots
