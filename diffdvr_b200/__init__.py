"""B200-native differentiable volume ray marcher -- drop-in for the hot path of DiffDVR.

Only what the path needs lives here:
  csrc/        hand-written CUDA (sm_100a) + the C-ABI (include/dvr_b200.h) -> libdvr_b200.so
  capi.py      ctypes mirror of the C-ABI (host plumbing; no CPU fallback)
  synthetic.py deterministic synthetic inputs (volumes, cameras, TFs) for tests and bench
  multigpu.py  view sharding + NCCL all-reduce of the volume / TF gradients
"""
from . import capi  # noqa: F401

__all__ = ["capi"]
