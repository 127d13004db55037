"""Drop-in for lzero/mcts/ctree/ctree_efficientzero/ez_tree.pyx over the CUDA forest (variant LZB_EZ:
value_prefix differencing and is_reset, ctree_efficientzero/lib/cnode.cpp:185-196,506-572)."""
from .._tree_api import MinMaxStatsList, NodeBase, ResultsWrapper, RootsBase, batch_traverse, set_tie_break  # noqa: F401
from .._tree_api import batch_backpropagate_ez as batch_backpropagate  # noqa: F401
from .._tree_api import batch_backpropagate_with_reuse_ez as batch_backpropagate_with_reuse  # noqa: F401
from .._tree_api import batch_traverse_with_reuse  # noqa: F401


class Roots(RootsBase):
    """ez_tree.pyx:26-60."""
    variant = "ez"


class Node(NodeBase):
    """ez_tree.pyx:70-81."""
    _roots_cls = Roots
