"""B200-native batched minimum-time trajectory optimiser for the Franka Panda (hot path of
AlbericDeLajarte/mpc_motion_planner behind a C ABI).  The compute lives in libmpcb.so (CUDA, sm_100a);
this package only binds it (ctypes) for the tests and the benchmark.  There is no CPU fallback."""
from .capi import (Account, Limits, MpcbError, OtgTraj, Settings, Solver, Stats, default_limits, default_settings, generate_problems,
                   guess_batch, lib_path, load_library, otg_at_time, otg_calculate, time_nodes)

__all__ = ["Account", "Limits", "MpcbError", "OtgTraj", "Settings", "Solver", "Stats", "default_limits", "default_settings",
           "generate_problems", "guess_batch", "lib_path", "load_library", "otg_at_time", "otg_calculate", "time_nodes"]
