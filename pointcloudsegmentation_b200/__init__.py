"""pointcloudsegmentation_b200 — B200-native (sm_100a) replacement for the tf_ops/ neighbor-operator
layer of liuyuan-pal/PointCloudSegmentation.

    from pointcloudsegmentation_b200 import neighbor_ops, pooling_ops          # raw op namespaces
    from pointcloudsegmentation_b200.graph_layers import graph_pool, ...        # L3 wrappers

Importing the package loads pointcloudsegmentation_b200/lib/libpcseg_ops.so (building it with nvcc if
absent).  There is no CPU or PyTorch fallback anywhere in the package.
"""
from . import _lib  # noqa: F401  (fail loudly at import if the CUDA library is unavailable)
from . import graph_layers, neighbor_ops, ops, pooling_ops  # noqa: F401
from ._lib import LIB_PATH, PcsError, launch_count  # noqa: F401

__version__ = "0.1.0"
