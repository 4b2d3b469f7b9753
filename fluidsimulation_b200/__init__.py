"""fluidsimulation_b200: B200-native (sm_100a CUDA) replacement for the reference CPU solver FluidSim3.

The product is the C-ABI shared library libfluidsim3_b200.so (include/fs3.h); `FluidSim3` here is the
host-side mirror of the reference class over that ABI.
"""
from .sim import (FluidSim3, FluidSim3Group, Fs3Error, Fs3Params, load_library, shard_plan, merge_particles, torch_allgather,  # noqa: F401
                  LIB_PATH, ABI_SYMBOLS, STAGE_NAMES, SHARD_HANDLE_BYTES, Fs3StateInfo, state_checksum, state_file_info,
                  state_file_write, state_file_read)
