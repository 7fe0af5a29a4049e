// Fused aggregate+apply, bulk-async (TMA) staged variant: declaration.
#pragma once

#include "ps_common.h"

namespace accel {

// Returns ACCEL_ERR_STATE (without setting an error message) when the shape is
// not covered by this variant; the caller then uses the vector-load variant.
int launch_apply_tma(int optimizer, int grad_dtype, const ApplyArgs &a, const TaskState &ts,
                     const Tuning &tune, cudaStream_t stream);

}  // namespace accel
