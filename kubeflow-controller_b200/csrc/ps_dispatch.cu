// Variant selection for the fused aggregate+apply launch. The variant is an
// explicit choice stored in the context (accel_ctx_set_kernel); AUTO is the
// shape measured fastest on B200 for the BASELINE configs.
#include "ps_apply_ldg.cuh"
#include "ps_apply_tma.cuh"

namespace accel {

template <typename G>
static int launch_opt(int optimizer, const ApplyArgs &a, bool wide, const LaunchShape &ls,
                      cudaStream_t stream) {
  switch (optimizer) {
    case ACCEL_OPT_SGD:
      return launch_apply_ldg<OPT_SGD, G>(a, a.n_workers, wide, ls, stream);
    case ACCEL_OPT_MOMENTUM:
      return launch_apply_ldg<OPT_MOMENTUM, G>(a, a.n_workers, wide, ls, stream);
    case ACCEL_OPT_ADAM:
      return launch_apply_ldg<OPT_ADAM, G>(a, a.n_workers, wide, ls, stream);
  }
  return fail(ACCEL_ERR_INVALID, "unknown optimizer %d", optimizer);
}

int launch_apply(int optimizer, int grad_dtype, const ApplyArgs &a, const TaskState &ts,
                 const Tuning &tune, cudaStream_t stream) {
  int kernel = tune.kernel;
  if (kernel == ACCEL_KERNEL_AUTO) {
    // 256-bit loads measured ~5% faster than 128-bit on the BASELINE configs; they
    // need 32-byte aligned buffers, which AUTO checks (an explicit LDG256 request
    // with misaligned buffers is an error instead, see build_args).
    uintptr_t bits = reinterpret_cast<uintptr_t>(a.var) | reinterpret_cast<uintptr_t>(a.slot0) |
                     reinterpret_cast<uintptr_t>(a.slot1);
    for (int w = 0; w < a.n_workers; ++w) bits |= reinterpret_cast<uintptr_t>(a.grad[w]);
    kernel = (bits & 31) ? ACCEL_KERNEL_LDG128 : ACCEL_KERNEL_LDG256;
  }
  if (kernel == ACCEL_KERNEL_TMA) {
    int rc = launch_apply_tma(optimizer, grad_dtype, a, ts, tune, stream);
    if (rc != ACCEL_ERR_STATE) return rc;  // ERR_STATE: shape not covered by the TMA variant
    kernel = ACCEL_KERNEL_LDG128;
  }
  LaunchShape ls;
  ls.kernel = kernel;
  ls.threads = tune.threads ? tune.threads : 256;
  ls.unroll = tune.unroll ? tune.unroll : 2;
  // grid <= 0: resident CTAs per SM from the occupancy calculator (-k caps it at k)
  ls.grid = tune.ctas_per_sm ? -tune.ctas_per_sm : 0;
  const bool wide = kernel == ACCEL_KERNEL_LDG256;
  return grad_dtype == ACCEL_BF16 ? launch_opt<bf16_t>(optimizer, a, wide, ls, stream)
                                  : launch_opt<float>(optimizer, a, wide, ls, stream);
}

}  // namespace accel
