// Instantiations of the fused aggregate+apply kernel: optimizer MOMENTUM, f32 gradients.
#include "ps_apply_inst.cuh"

ACCEL_INSTANTIATE_APPLY_LDG(accel::OPT_MOMENTUM, float)
