"""tiler_b200 — B200-native FrameTiling best-tile search + KModes assignment.

The product is tiler_b200/libtmann.so (hand-written sm_100a CUDA behind the
reference's ANN.dll C ABI, include/tmann.h).  This package is the thin host
mirror used by tests and bench.py.  Importing it requires the built library.
"""
from . import ann  # noqa: F401
from .ann import (ENGINE_AUTO, ENGINE_SIMT, ENGINE_TC, KM_ASM, KM_GENERIC, KDTree, KModesState,  # noqa: F401
                  kmodes_assign, kmodes_assign_binned, kmodes_farthest_first, kmodes_rows, kmodes_representatives,
                  kmodes_update_min_distance, psyv_queries, reindex, reset_stats,
                  comm_init_ms, get_devices, set_device, set_devices, set_engine, set_profiling, set_stream, stats, temporal_smoothing)
