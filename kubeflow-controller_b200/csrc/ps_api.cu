// The C ABI of include/accel_ps.h: context, argument checking, launches, and
// the host-buffer PS step. No torch types, no exceptions across the boundary.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "ps_common.h"

namespace accel {

// ---- errors ---------------------------------------------------------------
static thread_local char g_err[768] = "";

int fail(int status, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return status;
}

int fail_cuda(cudaError_t e, const char *what) {
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) in %s", static_cast<int>(e),
           cudaGetErrorString(e), what);
  cudaGetLastError();  // clear the sticky-free error so the next call starts clean
  return ACCEL_ERR_CUDA;
}

// ---- per-task state ---------------------------------------------------------
int task_state(accel_ctx *ctx, int task, TaskState **out) {
  if (!ctx) return fail(ACCEL_ERR_INVALID, "ctx is NULL");
  if (task < 0 || task >= ctx->n_ps)
    return fail(ACCEL_ERR_INVALID, "task %d out of range [0,%d)", task, ctx->n_ps);
  TaskState &ts = ctx->task[task];
  ACCEL_CUDA(cudaSetDevice(ctx->gpu[task]));
  if (!ts.ready) {
    ts.device = ctx->gpu[task];
    ACCEL_CUDA(cudaDeviceGetAttribute(&ts.sm_count, cudaDevAttrMultiProcessorCount, ts.device));
    int major = 0;
    ACCEL_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, ts.device));
    if (major != 10)
      return fail(ACCEL_ERR_CUDA,
                  "device %d has compute capability %d.x; this library holds sm_100a code only",
                  ts.device, major);
    // scratch of the sum-of-squares kernel: partials | ticket | local result
    const size_t bytes = sizeof(float) * kReducePartials + 256;
    void *p = nullptr;
    ACCEL_CUDA(cudaMalloc(&p, bytes));
    ACCEL_CUDA(cudaMemset(p, 0, bytes));
    ts.reduce_partials = static_cast<float *>(p);
    ts.reduce_ticket = reinterpret_cast<unsigned int *>(ts.reduce_partials + kReducePartials);
    ts.sumsq_local = ts.reduce_partials + kReducePartials + 32;
    ts.ready = true;
  }
  *out = &ts;
  return ACCEL_OK;
}

static void task_release(TaskState &ts) {
  if (!ts.ready) return;
  cudaSetDevice(ts.device);
  cudaFree(ts.reduce_partials);
  for (int i = 0; i < 2; ++i) {
    if (ts.stage_buf[i]) cudaFree(ts.stage_buf[i]);
    if (ts.ev_in[i]) cudaEventDestroy(ts.ev_in[i]);
    if (ts.ev_free[i]) cudaEventDestroy(ts.ev_free[i]);
    if (ts.ev_done[i]) cudaEventDestroy(ts.ev_done[i]);
  }
  if (ts.copy_in) cudaStreamDestroy(ts.copy_in);
  if (ts.compute) cudaStreamDestroy(ts.compute);
  if (ts.copy_out) cudaStreamDestroy(ts.copy_out);
  if (ts.shard_aux) cudaStreamDestroy(ts.shard_aux);
  if (ts.shard_ev_push) cudaEventDestroy(ts.shard_ev_push);
  if (ts.shard_ev_done) cudaEventDestroy(ts.shard_ev_done);
  if (ts.host_step_grads) cudaFree(ts.host_step_grads);
  for (int i = 0; i < TaskState::kHostSlices; ++i) {
    if (ts.ev_slice_in[i]) cudaEventDestroy(ts.ev_slice_in[i]);
    if (ts.ev_slice_done[i]) cudaEventDestroy(ts.ev_slice_done[i]);
  }
  ts = TaskState();
}

// ---- argument block ---------------------------------------------------------
static int build_args(int optimizer, float *var, float *slot0, float *slot1,
                      const void *const *grads, int n_workers, int grad_dtype, uint64_t count,
                      const accel_hparams *hp, const float *sumsq, int kernel, ApplyArgs *a) {
  if (!hp) return fail(ACCEL_ERR_INVALID, "hparams is NULL");
  const bool reduce_only = optimizer < 0;  // sum-of-squares: gradients only, no state
  if (!var && !reduce_only) return fail(ACCEL_ERR_INVALID, "var is NULL");
  if (!grads) return fail(ACCEL_ERR_INVALID, "grads is NULL");
  if (n_workers < 1 || n_workers > ACCEL_MAX_WORKERS)
    return fail(ACCEL_ERR_INVALID, "n_workers %d out of range [1,%d]", n_workers,
                ACCEL_MAX_WORKERS);
  if (grad_dtype != ACCEL_F32 && grad_dtype != ACCEL_BF16)
    return fail(ACCEL_ERR_INVALID, "unknown grad_dtype %d", grad_dtype);
  if (hp->aggregate != ACCEL_AGG_SUM && hp->aggregate != ACCEL_AGG_MEAN)
    return fail(ACCEL_ERR_INVALID, "unknown aggregate %d", hp->aggregate);
  if (optimizer >= ACCEL_OPT_MOMENTUM && !slot0)
    return fail(ACCEL_ERR_INVALID, "optimizer slot 0 (accum / m) is NULL");
  if (optimizer == ACCEL_OPT_ADAM && !slot1)
    return fail(ACCEL_ERR_INVALID, "optimizer slot 1 (v) is NULL");
  const uintptr_t mask = (kernel == ACCEL_KERNEL_LDG256) ? 31u : 15u;
  auto misaligned = [mask](const void *p) { return (reinterpret_cast<uintptr_t>(p) & mask) != 0; };
  if ((var && misaligned(var)) || (slot0 && misaligned(slot0)) || (slot1 && misaligned(slot1)))
    return fail(ACCEL_ERR_ALIGN, "var/slot pointers must be %u-byte aligned",
                static_cast<unsigned>(mask + 1));
  memset(a, 0, sizeof(*a));
  for (int w = 0; w < n_workers; ++w) {
    if (!grads[w]) return fail(ACCEL_ERR_INVALID, "grads[%d] is NULL", w);
    if (misaligned(grads[w]))
      return fail(ACCEL_ERR_ALIGN, "grads[%d] must be %u-byte aligned", w,
                  static_cast<unsigned>(mask + 1));
    a->grad[w] = grads[w];
  }
  a->var = var;
  a->slot0 = slot0;
  a->slot1 = slot1;
  a->sumsq = sumsq;
  a->count = count;
  a->n_workers = n_workers;
  a->use_nesterov = hp->use_nesterov ? 1 : 0;
  a->lr = hp->lr;
  a->momentum = hp->momentum;
  a->beta1 = hp->beta1;
  a->one_minus_beta1 = 1.0f - hp->beta1;
  a->one_minus_beta2 = 1.0f - hp->beta2;
  a->epsilon = hp->epsilon;
  a->alpha = 0.0f;
  if (optimizer == ACCEL_OPT_ADAM) {
    // TF-1.4 ApplyAdamNonCuda: alpha = lr * sqrt(1 - beta2_power) / (1 - beta1_power), all fp32
    const float num = hp->lr * sqrtf(1.0f - hp->beta2_power);
    a->alpha = num / (1.0f - hp->beta1_power);
  }
  a->mean_div = hp->aggregate == ACCEL_AGG_MEAN ? static_cast<float>(n_workers) : 0.0f;
  a->clip_norm = hp->clip_norm > 0.0f ? hp->clip_norm : 0.0f;
  return ACCEL_OK;
}

static int apply_common(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                        float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                        uint64_t count, const accel_hparams *hp, const float *sumsq_dev,
                        accel_stream_t stream, const float *dev_powers = nullptr) {
  TaskState *ts = nullptr;
  int rc = task_state(ctx, task, &ts);
  if (rc) return rc;
  ApplyArgs a;
  rc = build_args(optimizer, var, slot0, slot1, grads, n_workers, grad_dtype, count, hp, sumsq_dev,
                  ctx->tuning.kernel, &a);
  if (rc) return rc;
  if (count == 0) return ACCEL_OK;
  a.dev_powers = optimizer == ACCEL_OPT_ADAM ? dev_powers : nullptr;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a.clip_norm > 0.0f && a.sumsq == nullptr) {
    rc = launch_sumsq(a, grad_dtype, *ts, ts->sumsq_local, s);
    if (rc) return rc;
    a.sumsq = ts->sumsq_local;
  }
  return launch_apply(optimizer, grad_dtype, a, *ts, ctx->tuning, s);
}

int ensure_host_streams(TaskState &ts) {
  if (!ts.copy_in) {
    ACCEL_CUDA(cudaStreamCreateWithFlags(&ts.copy_in, cudaStreamNonBlocking));
    ACCEL_CUDA(cudaStreamCreateWithFlags(&ts.compute, cudaStreamNonBlocking));
    ACCEL_CUDA(cudaStreamCreateWithFlags(&ts.copy_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      ACCEL_CUDA(cudaEventCreateWithFlags(&ts.ev_in[i], cudaEventDisableTiming));
      ACCEL_CUDA(cudaEventCreateWithFlags(&ts.ev_free[i], cudaEventDisableTiming));
      ACCEL_CUDA(cudaEventCreateWithFlags(&ts.ev_done[i], cudaEventDisableTiming));
    }
  }
  return ACCEL_OK;
}

static int ensure_host_step(TaskState &ts, size_t stage_bytes) {
  int rc = ensure_host_streams(ts);
  if (rc) return rc;
  if (ts.stage_bytes < stage_bytes) {
    for (int i = 0; i < 2; ++i) {
      if (ts.stage_buf[i]) ACCEL_CUDA(cudaFree(ts.stage_buf[i]));
      ts.stage_buf[i] = nullptr;
    }
    ts.stage_bytes = 0;
    for (int i = 0; i < 2; ++i) ACCEL_CUDA(cudaMalloc(&ts.stage_buf[i], stage_bytes));
    ts.stage_bytes = stage_bytes;
  }
  return ACCEL_OK;
}

}  // namespace accel

using namespace accel;

extern "C" {

int accel_version(void) { return ACCEL_PS_VERSION; }

const char *accel_last_error(void) { return g_err; }

int accel_ctx_create(const int *gpu_ordinals, int n_ps, accel_ctx **out) {
  if (!out) return fail(ACCEL_ERR_INVALID, "out is NULL");
  *out = nullptr;
  if (n_ps < 1 || n_ps > ACCEL_MAX_PS)
    return fail(ACCEL_ERR_INVALID, "n_ps %d out of range [1,%d]", n_ps, ACCEL_MAX_PS);
  accel_ctx *c = new (std::nothrow) accel_ctx();
  if (!c) return fail(ACCEL_ERR_NOMEM, "out of host memory");
  c->n_ps = n_ps;
  const char *tr = getenv("ACCEL_STEP_TRACE");   // development knob, see DESIGN.md
  c->tuning.trace = tr && tr[0] == '1';
  const char *l2 = getenv("ACCEL_STEP_L2HINT");
  c->tuning.l2hint = l2 && l2[0] == '1';
  for (int i = 0; i < n_ps; ++i) {
    c->gpu[i] = gpu_ordinals ? gpu_ordinals[i] : i;
    if (c->gpu[i] < 0) {
      const int bad = c->gpu[i];
      delete c;
      return fail(ACCEL_ERR_INVALID, "gpu_ordinals[%d] = %d is negative", i, bad);
    }
  }
  *out = c;
  return ACCEL_OK;
}

int accel_ctx_destroy(accel_ctx *ctx) {
  if (!ctx) return ACCEL_OK;
  for (int i = 0; i < ctx->n_ps; ++i) task_release(ctx->task[i]);
  delete ctx;
  return ACCEL_OK;
}

int accel_ctx_gpu_of(const accel_ctx *ctx, int task) {
  if (!ctx) return fail(ACCEL_ERR_INVALID, "ctx is NULL");
  if (task < 0 || task >= ctx->n_ps)
    return fail(ACCEL_ERR_INVALID, "task %d out of range [0,%d)", task, ctx->n_ps);
  return ctx->gpu[task];
}

int accel_ctx_set_kernel(accel_ctx *ctx, int kernel, int threads, int ctas_per_sm, int unroll) {
  if (!ctx) return fail(ACCEL_ERR_INVALID, "ctx is NULL");
  if (kernel < ACCEL_KERNEL_AUTO || kernel > ACCEL_KERNEL_TMA)
    return fail(ACCEL_ERR_INVALID, "unknown kernel %d", kernel);
  if (threads != 0 && (threads < 32 || threads > 512 || threads % 32))
    return fail(ACCEL_ERR_INVALID, "threads %d must be a multiple of 32 in [32,512]", threads);
  if (ctas_per_sm < 0 || ctas_per_sm > 32)
    return fail(ACCEL_ERR_INVALID, "ctas_per_sm %d out of range [0,32]", ctas_per_sm);
  if (unroll != 0 && unroll != 1 && unroll != 2 && unroll != 4)
    return fail(ACCEL_ERR_INVALID, "unroll %d must be 1, 2 or 4", unroll);
  ctx->tuning.kernel = kernel;
  ctx->tuning.threads = threads;
  ctx->tuning.ctas_per_sm = ctas_per_sm;
  ctx->tuning.unroll = unroll;
  return ACCEL_OK;
}

int accel_apply_sgd(accel_ctx *ctx, int task, float *var, const void *const *grads, int n_workers,
                    int grad_dtype, uint64_t count, const accel_hparams *hp,
                    const float *grad_sumsq_dev, accel_stream_t stream) {
  return apply_common(ctx, task, ACCEL_OPT_SGD, var, nullptr, nullptr, grads, n_workers,
                      grad_dtype, count, hp, grad_sumsq_dev, stream);
}

int accel_apply_momentum(accel_ctx *ctx, int task, float *var, float *accum,
                         const void *const *grads, int n_workers, int grad_dtype, uint64_t count,
                         const accel_hparams *hp, const float *grad_sumsq_dev,
                         accel_stream_t stream) {
  return apply_common(ctx, task, ACCEL_OPT_MOMENTUM, var, accum, nullptr, grads, n_workers,
                      grad_dtype, count, hp, grad_sumsq_dev, stream);
}

int accel_apply_adam(accel_ctx *ctx, int task, float *var, float *m, float *v,
                     const void *const *grads, int n_workers, int grad_dtype, uint64_t count,
                     const accel_hparams *hp, const float *grad_sumsq_dev, accel_stream_t stream) {
  return apply_common(ctx, task, ACCEL_OPT_ADAM, var, m, v, grads, n_workers, grad_dtype, count,
                      hp, grad_sumsq_dev, stream);
}

int accel_opt_state_init(void *dev_state, const accel_hparams *hp, uint64_t global_step,
                         accel_stream_t stream) {
  if (!dev_state || !hp) return fail(ACCEL_ERR_INVALID, "dev_state/hparams is NULL");
  if (reinterpret_cast<uintptr_t>(dev_state) & 15)
    return fail(ACCEL_ERR_ALIGN, "dev_state must be 16-byte aligned");
  return launch_opt_init(dev_state, hp->beta1_power, hp->beta2_power, global_step,
                         static_cast<cudaStream_t>(stream));
}

int accel_opt_state_advance(void *dev_state, const accel_hparams *hp, int is_adam,
                            accel_stream_t stream) {
  if (!dev_state || !hp) return fail(ACCEL_ERR_INVALID, "dev_state/hparams is NULL");
  if (reinterpret_cast<uintptr_t>(dev_state) & 15)
    return fail(ACCEL_ERR_ALIGN, "dev_state must be 16-byte aligned");
  return launch_opt_advance(dev_state, hp->beta1, hp->beta2, is_adam != 0,
                            static_cast<cudaStream_t>(stream));
}

int accel_apply_dev_state(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                          float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                          uint64_t count, const accel_hparams *hp, const float *grad_sumsq_dev,
                          void *dev_state, accel_stream_t stream) {
  return accel_apply_dev_state_ex(ctx, task, optimizer, var, slot0, slot1, grads, n_workers,
                                  grad_dtype, count, hp, grad_sumsq_dev, dev_state, /*advance=*/1,
                                  stream);
}

int accel_apply_dev_state_ex(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                             float *slot1, const void *const *grads, int n_workers,
                             int grad_dtype, uint64_t count, const accel_hparams *hp,
                             const float *grad_sumsq_dev, void *dev_state, int advance,
                             accel_stream_t stream) {
  if (optimizer < ACCEL_OPT_SGD || optimizer > ACCEL_OPT_ADAM)
    return fail(ACCEL_ERR_INVALID, "unknown optimizer %d", optimizer);
  if (!dev_state) return fail(ACCEL_ERR_INVALID, "dev_state is NULL");
  if (reinterpret_cast<uintptr_t>(dev_state) & 15)
    return fail(ACCEL_ERR_ALIGN, "dev_state must be 16-byte aligned");
  if (!hp) return fail(ACCEL_ERR_INVALID, "hparams is NULL");
  int rc = apply_common(ctx, task, optimizer, var, optimizer >= ACCEL_OPT_MOMENTUM ? slot0 : nullptr,
                        optimizer == ACCEL_OPT_ADAM ? slot1 : nullptr, grads, n_workers,
                        grad_dtype, count, hp, grad_sumsq_dev, stream,
                        static_cast<const float *>(dev_state));
  if (rc || !advance) return rc;
  return launch_opt_advance(dev_state, hp->beta1, hp->beta2, optimizer == ACCEL_OPT_ADAM,
                            static_cast<cudaStream_t>(stream));
}

int accel_adam_advance(accel_hparams *hp) {
  if (!hp) return fail(ACCEL_ERR_INVALID, "hparams is NULL");
  hp->beta1_power = hp->beta1_power * hp->beta1;
  hp->beta2_power = hp->beta2_power * hp->beta2;
  return ACCEL_OK;
}

int accel_grad_sumsq(accel_ctx *ctx, int task, const void *const *grads, int n_workers,
                     int grad_dtype, uint64_t count, int aggregate, float *out_sumsq_dev,
                     accel_stream_t stream) {
  TaskState *ts = nullptr;
  int rc = task_state(ctx, task, &ts);
  if (rc) return rc;
  if (!out_sumsq_dev) return fail(ACCEL_ERR_INVALID, "out_sumsq_dev is NULL");
  accel_hparams hp;
  memset(&hp, 0, sizeof(hp));
  hp.aggregate = aggregate;
  ApplyArgs a;
  rc = build_args(/*optimizer=*/-1, nullptr, nullptr, nullptr, grads, n_workers, grad_dtype, count,
                  &hp, nullptr, ACCEL_KERNEL_LDG128, &a);
  if (rc) return rc;
  if (count == 0) {
    ACCEL_CUDA(cudaMemsetAsync(out_sumsq_dev, 0, sizeof(float), static_cast<cudaStream_t>(stream)));
    return ACCEL_OK;
  }
  return launch_sumsq(a, grad_dtype, *ts, out_sumsq_dev, static_cast<cudaStream_t>(stream));
}

// ---- the host-buffer step ---------------------------------------------------
int accel_ps_step_host(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                       float *slot1, const void *const *host_grads, int n_workers, int grad_dtype,
                       uint64_t count, const accel_hparams *hp, float *host_var_out) {
  TaskState *ts = nullptr;
  int rc = task_state(ctx, task, &ts);
  if (rc) return rc;
  if (optimizer < ACCEL_OPT_SGD || optimizer > ACCEL_OPT_ADAM)
    return fail(ACCEL_ERR_INVALID, "unknown optimizer %d", optimizer);
  if (!hp) return fail(ACCEL_ERR_INVALID, "hparams is NULL");
  if (hp->clip_norm > 0.0f)
    return fail(ACCEL_ERR_INVALID,
                "accel_ps_step_host streams the gradient in chunks and cannot clip by global "
                "norm; clip_norm must be 0");
  if (!host_grads) return fail(ACCEL_ERR_INVALID, "host_grads is NULL");
  if (n_workers < 1 || n_workers > ACCEL_MAX_WORKERS)
    return fail(ACCEL_ERR_INVALID, "n_workers %d out of range [1,%d]", n_workers,
                ACCEL_MAX_WORKERS);
  if (grad_dtype != ACCEL_F32 && grad_dtype != ACCEL_BF16)
    return fail(ACCEL_ERR_INVALID, "unknown grad_dtype %d", grad_dtype);
  if (count == 0) return ACCEL_OK;
  const size_t gsize = grad_dtype == ACCEL_BF16 ? 2 : 4;
  // chunk: 4 Mi elements per worker (16 MiB fp32) keeps each H2D copy well past
  // the PCIe latency knee while the pipeline has >= 20 chunks at BASELINE size
  const uint64_t chunk = std::min<uint64_t>(count, uint64_t{4} << 20);
  const size_t per_worker = (chunk * gsize + 255) & ~size_t{255};
  rc = ensure_host_step(*ts, per_worker * n_workers);
  if (rc) return rc;

  for (int w = 0; w < n_workers; ++w)
    if (!host_grads[w]) return fail(ACCEL_ERR_INVALID, "host_grads[%d] is NULL", w);
  const uint64_t n_chunks = (count + chunk - 1) / chunk;
  // the pipeline proper; any failure falls through to the drain below so that no copy or
  // kernel is still in flight on the caller's host buffers when the call returns
  auto enqueue = [&]() -> int {
    for (uint64_t k = 0; k < n_chunks; ++k) {
      const int b = static_cast<int>(k & 1);
      const uint64_t off = k * chunk;
      const uint64_t n = std::min<uint64_t>(chunk, count - off);
      // copy-in: wait until the kernel that last read this staging buffer is done
      if (k >= 2) ACCEL_CUDA(cudaStreamWaitEvent(ts->copy_in, ts->ev_free[b], 0));
      const void *dev_grads[ACCEL_MAX_WORKERS];
      for (int w = 0; w < n_workers; ++w) {
        char *dst = static_cast<char *>(ts->stage_buf[b]) + per_worker * w;
        const char *src = static_cast<const char *>(host_grads[w]) + off * gsize;
        ACCEL_CUDA(cudaMemcpyAsync(dst, src, n * gsize, cudaMemcpyHostToDevice, ts->copy_in));
        dev_grads[w] = dst;
      }
      ACCEL_CUDA(cudaEventRecord(ts->ev_in[b], ts->copy_in));
      // compute
      ACCEL_CUDA(cudaStreamWaitEvent(ts->compute, ts->ev_in[b], 0));
      int rc2 = apply_common(ctx, task, optimizer, var + off, slot0 ? slot0 + off : nullptr,
                             slot1 ? slot1 + off : nullptr, dev_grads, n_workers, grad_dtype, n, hp,
                             nullptr, ts->compute);
      if (rc2) return rc2;
      ACCEL_CUDA(cudaEventRecord(ts->ev_free[b], ts->compute));
      // copy-out: the pull of the updated chunk
      if (host_var_out) {
        ACCEL_CUDA(cudaEventRecord(ts->ev_done[b], ts->compute));
        ACCEL_CUDA(cudaStreamWaitEvent(ts->copy_out, ts->ev_done[b], 0));
        ACCEL_CUDA(cudaMemcpyAsync(host_var_out + off, var + off, n * sizeof(float),
                                   cudaMemcpyDeviceToHost, ts->copy_out));
      }
    }
    return ACCEL_OK;
  };
  rc = enqueue();
  // drain all three streams on success AND on failure
  const cudaError_t e1 = cudaStreamSynchronize(ts->copy_in);
  const cudaError_t e2 = cudaStreamSynchronize(ts->compute);
  const cudaError_t e3 = cudaStreamSynchronize(ts->copy_out);
  if (rc) return rc;   // g_err holds the first failure's message
  if (e1 != cudaSuccess) return fail_cuda(e1, "cudaStreamSynchronize(copy_in)");
  if (e2 != cudaSuccess) return fail_cuda(e2, "cudaStreamSynchronize(compute)");
  if (e3 != cudaSuccess) return fail_cuda(e3, "cudaStreamSynchronize(copy_out)");
  return ACCEL_OK;
}

int accel_host_register(void *host_ptr, size_t bytes) {
  if (!host_ptr || !bytes) return fail(ACCEL_ERR_INVALID, "host_ptr/bytes is NULL/0");
  ACCEL_CUDA(cudaHostRegister(host_ptr, bytes, cudaHostRegisterPortable));
  return ACCEL_OK;
}

int accel_host_unregister(void *host_ptr) {
  if (!host_ptr) return fail(ACCEL_ERR_INVALID, "host_ptr is NULL");
  ACCEL_CUDA(cudaHostUnregister(host_ptr));
  return ACCEL_OK;
}

int accel_ipc_export(void *dev_ptr, accel_ipc_handle *out) {
  if (!dev_ptr || !out) return fail(ACCEL_ERR_INVALID, "dev_ptr/out is NULL");
  static_assert(sizeof(cudaIpcMemHandle_t) <= ACCEL_IPC_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  ACCEL_CUDA(cudaIpcGetMemHandle(&h, dev_ptr));
  memset(out, 0, sizeof(*out));
  memcpy(out->bytes, &h, sizeof(h));
  return ACCEL_OK;
}

int accel_ipc_open(const accel_ipc_handle *h, void **out_dev_ptr) {
  if (!h || !out_dev_ptr) return fail(ACCEL_ERR_INVALID, "handle/out is NULL");
  cudaIpcMemHandle_t ch;
  memcpy(&ch, h->bytes, sizeof(ch));
  ACCEL_CUDA(cudaIpcOpenMemHandle(out_dev_ptr, ch, cudaIpcMemLazyEnablePeerAccess));
  return ACCEL_OK;
}

int accel_ipc_close(void *dev_ptr) {
  if (!dev_ptr) return fail(ACCEL_ERR_INVALID, "dev_ptr is NULL");
  ACCEL_CUDA(cudaIpcCloseMemHandle(dev_ptr));
  return ACCEL_OK;
}

}  // extern "C"
