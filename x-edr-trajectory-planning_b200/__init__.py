"""B200-native batched path-timing engine (drop-in for the hot path of
theteamatx/x-edr-trajectory-planning).

The product is ``lib/libtpb.so`` (hand-written sm_100a CUDA kernels behind the C
ABI of ``include/tpb.h``) plus the host-side mirrors of the reference's classes:
C++ in ``host/`` (same names as the reference) and this thin ctypes layer used by
the tests and ``bench.py``.  There is no CPU fallback: if the library is missing
or no CUDA device is present, construction fails loudly.

The directory name contains hyphens, so import it through
``__graft_entry__.load_package()`` (module name ``x_edr_trajectory_planning_b200``).
"""
from .binding import (  # noqa: F401
    ARR, PATH_OK, Engine, TpbError, lib, lib_path, rescale_to_stop_device, rescale_to_stop_host,
    spline_eval_device, spline_eval_host, quat_spline_eval_device, quat_spline_eval_host,
    TrajectoryBufferBatch, TrajectoryBufferView, SKIP_INDEX,
    time_from_sec_ns, time_to_sec, measure_fp64_fma_peak, resample_uniform_noalloc, MultiEngine, shard_range, HostBatch,
)
from . import workloads  # noqa: F401
