#include "ps_apply_tma.cuh"

namespace accel {

int launch_apply_tma(int, int, const ApplyArgs &, const TaskState &, const Tuning &, cudaStream_t) {
  return ACCEL_ERR_STATE;
}

}  // namespace accel
