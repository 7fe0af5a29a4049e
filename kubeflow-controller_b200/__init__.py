"""B200-native parameter-server step for the TFJobs of gaocegege/kubeflow-controller.

The product is libaccel_ps.so (csrc/, C ABI in include/accel_ps.h): hand-written
sm_100a CUDA kernels for the fused N-way gradient aggregation + optimizer apply,
and the placement contract of pkg/tensorflow. This package is the Python harness
over that ABI (ctypes); the Go wrapper a maintainer would use is go/pkg/accel.
"""
from . import accel, placement  # noqa: F401
from .accel import (AGG_MEAN, AGG_SUM, BF16, F32, KERNEL_AUTO, KERNEL_LDG128, KERNEL_LDG256,  # noqa: F401
                    KERNEL_TMA, OPT_ADAM, OPT_MOMENTUM, OPT_SGD, AccelError, Context, HParams,
                    load_library, plan_flat, plan_round_robin)
from .ps import ParameterServer  # noqa: F401

__all__ = ["accel", "placement", "Context", "HParams", "ParameterServer", "AccelError",
           "load_library", "plan_flat", "plan_round_robin"]
