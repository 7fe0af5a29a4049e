// Instantiations of the fused aggregate+apply kernel: optimizer ADAM, bf16 gradients.
#include "ps_apply_inst.cuh"

ACCEL_INSTANTIATE_APPLY_LDG(accel::OPT_ADAM, accel::bf16_t)
