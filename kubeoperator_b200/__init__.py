"""korollup — B200-native (sm_100a) monitoring rollup for KubeOperator.

Only the Day-2 monitoring rollup of core/apps/kubeops_api is here: the C-ABI CUDA library
``libkorollup.so`` (csrc/, include/korollup.h), its ctypes shim (korollup.py) and the host-side mirror of the
reference's rollup interface (cluster_monitor.py, fleet.py).  Importing this package never touches CUDA.
"""
from .korollup import KoRollupError, Rollup, load_library  # noqa: F401

__version__ = "0.1.0"
