// Internal declarations shared by the translation units of libaccel_ps.so.
// Nothing here is part of the ABI; the ABI is include/accel_ps.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "accel_ps.h"

namespace accel {

// ---- error plumbing -------------------------------------------------------
int fail(int status, const char *fmt, ...) __attribute__((format(printf, 2, 3)));
int fail_cuda(cudaError_t e, const char *what);
#define ACCEL_CUDA(call)                                   \
  do {                                                     \
    cudaError_t e__ = (call);                              \
    if (e__ != cudaSuccess) return ::accel::fail_cuda(e__, #call); \
  } while (0)

// ---- launch configuration -------------------------------------------------
struct Tuning {
  int kernel = ACCEL_KERNEL_AUTO;
  int threads = 0;
  int ctas_per_sm = 0;
  int unroll = 0;
  int reducer_pct = 0;  // in-switch step: share of CTAs that reduce+apply+pull (0 = default)
  int coop = 1;         // one-kernel step: cooperative launch (0: plain launch for grids <= SMs)
  int trace = 0;        // development: the in-switch step stamps %globaltimer into flag words 128..255
  int l2hint = 0;       // development: L2 policies on the summers' loads and stores (ACCEL_STEP_L2HINT)
};

constexpr int kReducePartials = 4096;  // upper bound on CTAs of the sum-of-squares grid

// Per-task (= per-GPU) state owned by the context. Created lazily on first use
// so that a one-process-per-GPU deployment never opens a context on a peer GPU.
struct TaskState {
  bool ready = false;
  int device = -1;
  int sm_count = 0;
  // scratch for accel_grad_sumsq: [kReducePartials] CTA partials, then one
  // ticket word, then one float result used when the apply computes the norm.
  float *reduce_partials = nullptr;
  unsigned int *reduce_ticket = nullptr;
  float *sumsq_local = nullptr;
  // host-buffer step: staging buffers and streams
  cudaStream_t copy_in = nullptr, compute = nullptr, copy_out = nullptr;
  void *stage_buf[2] = {nullptr, nullptr};
  size_t stage_bytes = 0;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};
  cudaEvent_t ev_done[2] = {nullptr, nullptr};
  // sharded host-buffer step: full-arena device copies of the local workers' gradients
  void *host_step_grads = nullptr;
  size_t host_step_grad_bytes = 0;
  static constexpr int kHostSlices = 64;
  cudaEvent_t ev_slice_in[kHostSlices] = {}, ev_slice_done[kHostSlices] = {};
  unsigned int shard_epoch = 0;  // barrier generation of the multi-GPU step
  cudaStream_t shard_aux = nullptr;  // apply_pull of slice k runs here beside push of slice k+1
  cudaEvent_t shard_ev_push = nullptr, shard_ev_done = nullptr;
};

}  // namespace accel

struct accel_ctx {
  int n_ps;
  int gpu[ACCEL_MAX_PS];
  accel::Tuning tuning;
  accel::TaskState task[ACCEL_MAX_PS];
};

namespace accel {

int task_state(accel_ctx *ctx, int task, TaskState **out);  // sets the device
int ensure_host_streams(TaskState &ts);  // copy_in / compute / copy_out + their events

// ---- kernel parameter blocks (passed by value as __grid_constant__) --------
struct ApplyArgs {
  float *var;
  float *slot0;
  float *slot1;
  const void *grad[ACCEL_MAX_WORKERS];
  const float *sumsq;  // device scalar for clipping, or nullptr
  const float *dev_powers;  // device {beta1_power, beta2_power} (Adam), or nullptr: use `alpha`
  uint64_t count;
  int n_workers;
  int use_nesterov;
  float lr;
  float momentum;
  float beta1;
  float one_minus_beta1;
  float one_minus_beta2;
  float alpha;     // Adam: lr*sqrt(1-b2^t)/(1-b1^t)
  float epsilon;
  float mean_div;  // float(N) for MEAN, 0 for SUM
  float clip_norm;
};

struct LaunchShape {
  int kernel, threads, grid, unroll;
};

int launch_apply(int optimizer, int grad_dtype, const ApplyArgs &a, const TaskState &ts,
                 const Tuning &tune, cudaStream_t stream);
int launch_sumsq(const ApplyArgs &a, int grad_dtype, const TaskState &ts, float *out,
                 cudaStream_t stream);
int launch_opt_advance(void *dev_state, float beta1, float beta2, int is_adam, cudaStream_t stream);
int launch_opt_init(void *dev_state, float beta1_power, float beta2_power, uint64_t global_step,
                    cudaStream_t stream);

}  // namespace accel
