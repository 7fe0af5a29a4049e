// Instantiations of the fused aggregate+apply kernel: optimizer SGD, bf16 gradients.
#include "ps_apply_inst.cuh"

ACCEL_INSTANTIATE_APPLY_LDG(accel::OPT_SGD, accel::bf16_t)
