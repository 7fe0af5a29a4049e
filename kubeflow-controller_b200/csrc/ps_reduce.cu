// Sum of squares of the aggregated gradient, for tf.clip_by_global_norm
// (TF-1.4 clip_ops.py global_norm — not in /root/reference).
//
// One read pass over the N worker gradients (N*sizeof(g) bytes/element). Each
// thread accumulates G^2 in fp32 over a grid-strided set of vectors; the CTA
// combines with warp shuffles and one shared-memory hop; CTA partials go to a
// scratch array in global memory; the last CTA to finish (ticket counter) adds
// the partials in index order in fp64 and writes one fp32 result. For a given
// launch shape the result is bit-reproducible run to run.
#include <mutex>

#include "ps_math.cuh"

namespace accel {

constexpr int kReduceThreads = 512;

template <typename G, int NW>
__global__ void __launch_bounds__(kReduceThreads)
    ps_sumsq(const __grid_constant__ ApplyArgs a, float *__restrict__ partials,
             unsigned int *__restrict__ ticket, float *__restrict__ out) {
  constexpr int VE = Item<G, false>::VE;
  const bool has_div = a.mean_div != 0.0f;
  const uint64_t n_items = a.count / VE;
  const int n = NW > 0 ? NW : a.n_workers;

  float ss = 0.0f;
  for (uint64_t it = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; it < n_items;
       it += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    float acc[VE];
    if constexpr (NW > 0) {
      float g[NW][VE];
#pragma unroll
      for (int w = 0; w < NW; ++w) load_grad_item<G, false>(a.grad[w], it, g[w]);
#pragma unroll
      for (int e = 0; e < VE; ++e) {
        acc[e] = g[0][e];
#pragma unroll
        for (int w = 1; w < NW; ++w) acc[e] = __fadd_rn(acc[e], g[w][e]);
      }
    } else {
      load_grad_item<G, false>(a.grad[0], it, acc);
      for (int w = 1; w < n; ++w) {
        float g[VE];
        load_grad_item<G, false>(a.grad[w], it, g);
#pragma unroll
        for (int e = 0; e < VE; ++e) acc[e] = __fadd_rn(acc[e], g[e]);
      }
    }
#pragma unroll
    for (int e = 0; e < VE; ++e) {
      float x = has_div ? __fdiv_rn(acc[e], a.mean_div) : acc[e];
      ss = fmaf(x, x, ss);
    }
  }
  // scalar tail
  const uint64_t tail0 = n_items * VE;
  if (blockIdx.x == 0 && tail0 + threadIdx.x < a.count) {
    const uint64_t i = tail0 + threadIdx.x;
    float sum = load_grad_scalar<G>(a.grad[0], i);
    for (int w = 1; w < n; ++w) sum = __fadd_rn(sum, load_grad_scalar<G>(a.grad[w], i));
    if (has_div) sum = __fdiv_rn(sum, a.mean_div);
    ss = fmaf(sum, sum, ss);
  }

  // warp shuffle tree, then one shared-memory hop across the warps of the CTA
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  __shared__ float warp_sum[kReduceThreads / 32];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_sum[warp] = ss;
  __syncthreads();
  if (warp == 0) {
    float t = lane < (blockDim.x >> 5) ? warp_sum[lane] : 0.0f;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
    if (lane == 0) {
      partials[blockIdx.x] = t;
      __threadfence();
      unsigned int done = atomicAdd(ticket, 1u);
      is_last = (done == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (is_last) {
    // every partial is visible: the writers fenced before taking a ticket
    __threadfence();
    double t = 0.0;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x)
      t += static_cast<double>(__ldcg(partials + i));
    // fixed-shape tree over the CTA in fp64
    __shared__ double dsum[kReduceThreads / 32];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
    if (lane == 0) dsum[warp] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      double tot = 0.0;
      for (int w = 0; w < (blockDim.x >> 5); ++w) tot += dsum[w];
      out[0] = static_cast<float>(tot);
      *ticket = 0u;  // ready for the next launch on this stream
    }
  }
}

template <typename G>
static int launch_sumsq_g(const ApplyArgs &a, const TaskState &ts, float *out,
                          cudaStream_t stream) {
  constexpr int VE = Item<G, false>::VE;
  const uint64_t items = a.count / VE;
  uint64_t want = (items + kReduceThreads - 1) / kReduceThreads;
  int grid = ts.sm_count * 4;  // 4 x 512 threads = all 2048 thread slots of an SM
  if (grid > kReducePartials) grid = kReducePartials;
  if (static_cast<uint64_t>(grid) > want) grid = want ? static_cast<int>(want) : 1;
#define ACCEL_SUMSQ(NW) \
  ps_sumsq<G, NW><<<grid, kReduceThreads, 0, stream>>>(a, ts.reduce_partials, ts.reduce_ticket, out)
  switch (a.n_workers) {
    case 1: ACCEL_SUMSQ(1); break;
    case 2: ACCEL_SUMSQ(2); break;
    case 4: ACCEL_SUMSQ(4); break;
    case 8: ACCEL_SUMSQ(8); break;
    default: ACCEL_SUMSQ(0); break;
  }
#undef ACCEL_SUMSQ
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

int launch_sumsq(const ApplyArgs &a, int grad_dtype, const TaskState &ts, float *out,
                 cudaStream_t stream) {
  return grad_dtype == ACCEL_BF16 ? launch_sumsq_g<bf16_t>(a, ts, out, stream)
                                  : launch_sumsq_g<float>(a, ts, out, stream);
}

}  // namespace accel

// ---- optimizer state that advances on the device ------------------------------------------
// Layout of the 16-byte device block: float beta1_power, float beta2_power, uint64 global_step.
namespace accel {

__global__ void ps_opt_init(float *state, float b1p, float b2p, unsigned long long step) {
  state[0] = b1p;
  state[1] = b2p;
  *reinterpret_cast<unsigned long long *>(state + 2) = step;
}

// TF-1.4 AdamOptimizer._finish: beta_power *= beta (fp32); global_step += 1 per apply.
__global__ void ps_opt_advance(float *state, float beta1, float beta2, int is_adam) {
  if (is_adam) {
    state[0] = __fmul_rn(state[0], beta1);
    state[1] = __fmul_rn(state[1], beta2);
  }
  *reinterpret_cast<unsigned long long *>(state + 2) += 1ull;
}

int launch_opt_init(void *dev_state, float b1p, float b2p, uint64_t step, cudaStream_t stream) {
  ps_opt_init<<<1, 1, 0, stream>>>(static_cast<float *>(dev_state), b1p, b2p, step);
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

int launch_opt_advance(void *dev_state, float beta1, float beta2, int is_adam,
                       cudaStream_t stream) {
  ps_opt_advance<<<1, 1, 0, stream>>>(static_cast<float *>(dev_state), beta1, beta2, is_adam);
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

}  // namespace accel

// ---- the entry point that allocates nothing -----------------------------------------
//
// Same kernel, but its scratch (CTA partials + the ticket word) is this file-scope
// device array instead of memory the context allocated. Every GPU that loads the
// library carries these 16.25 KiB until the process exits.
namespace accel {

constexpr int kStaticMaxCtas = 4096;                    // one fp32 partial per CTA
constexpr uint64_t kStaticElemsPerCta = 65536;          // elements a CTA strides over
constexpr size_t kStaticTicketOffset = kStaticMaxCtas * sizeof(float);
static_assert(uint64_t{kStaticMaxCtas} * kStaticElemsPerCta == ACCEL_SUMSQ_STATIC_MAX_ELEMS,
              "header constant out of step with the scratch size");

static __device__ __align__(256) unsigned char g_sumsq_static_scratch[kStaticTicketOffset + 256];

// In host code `g_sumsq_static_scratch` is only a handle for the runtime, not an address a
// kernel can use; the device address is looked up once per device and kept here.
constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char *g_static_addr[kMaxDevices] = {};

static int static_scratch_for_current_device(unsigned char **out) {
  int dev = -1;
  ACCEL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(ACCEL_ERR_INVALID, "device ordinal %d out of range [0,%d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[dev]) {
    void *p = nullptr;
    ACCEL_CUDA(cudaGetSymbolAddress(&p, g_sumsq_static_scratch));
    g_static_addr[dev] = static_cast<unsigned char *>(p);
  }
  *out = g_static_addr[dev];
  return ACCEL_OK;
}

template <typename G>
static int launch_sumsq_static(const ApplyArgs &a, int grid, float *partials, unsigned int *ticket,
                               float *out, cudaStream_t stream) {
#define ACCEL_SUMSQ_STATIC(NW) \
  ps_sumsq<G, NW><<<grid, kReduceThreads, 0, stream>>>(a, partials, ticket, out)
  switch (a.n_workers) {
    case 1: ACCEL_SUMSQ_STATIC(1); break;
    case 2: ACCEL_SUMSQ_STATIC(2); break;
    case 4: ACCEL_SUMSQ_STATIC(4); break;
    case 8: ACCEL_SUMSQ_STATIC(8); break;
    default: ACCEL_SUMSQ_STATIC(0); break;
  }
#undef ACCEL_SUMSQ_STATIC
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

}  // namespace accel

extern "C" int accel_grad_sumsq_static(const void *const *grads, int n_workers, int grad_dtype,
                                       uint64_t count, int aggregate, float *out_sumsq_dev,
                                       accel_stream_t stream) {
  using namespace accel;
  if (!grads) return fail(ACCEL_ERR_INVALID, "grads is NULL");
  if (!out_sumsq_dev) return fail(ACCEL_ERR_INVALID, "out_sumsq_dev is NULL");
  if (n_workers < 1 || n_workers > ACCEL_MAX_WORKERS)
    return fail(ACCEL_ERR_INVALID, "n_workers %d out of range [1,%d]", n_workers,
                ACCEL_MAX_WORKERS);
  if (grad_dtype != ACCEL_F32 && grad_dtype != ACCEL_BF16)
    return fail(ACCEL_ERR_INVALID, "unknown grad_dtype %d", grad_dtype);
  if (aggregate != ACCEL_AGG_SUM && aggregate != ACCEL_AGG_MEAN)
    return fail(ACCEL_ERR_INVALID, "unknown aggregate %d", aggregate);
  if (count > ACCEL_SUMSQ_STATIC_MAX_ELEMS)
    return fail(ACCEL_ERR_INVALID,
                "count %llu is beyond the %llu elements the built-in scratch of "
                "accel_grad_sumsq_static covers; split the range (sums of squares add) or use "
                "accel_grad_sumsq",
                static_cast<unsigned long long>(count),
                static_cast<unsigned long long>(ACCEL_SUMSQ_STATIC_MAX_ELEMS));
  ApplyArgs a;
  memset(&a, 0, sizeof(a));
  for (int w = 0; w < n_workers; ++w) {
    if (!grads[w]) return fail(ACCEL_ERR_INVALID, "grads[%d] is NULL", w);
    if (reinterpret_cast<uintptr_t>(grads[w]) & 15)
      return fail(ACCEL_ERR_ALIGN, "grads[%d] must be 16-byte aligned", w);
    a.grad[w] = grads[w];
  }
  a.count = count;
  a.n_workers = n_workers;
  a.mean_div = aggregate == ACCEL_AGG_MEAN ? static_cast<float>(n_workers) : 0.0f;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (count == 0) {
    ACCEL_CUDA(cudaMemsetAsync(out_sumsq_dev, 0, sizeof(float), s));
    return ACCEL_OK;
  }
  unsigned char *scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  float *partials = reinterpret_cast<float *>(scratch);
  unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + kStaticTicketOffset);
  // The array still holds whatever the previous call left. Partials need no clearing (each
  // CTA of this launch overwrites its own before the last CTA reads them); the ticket must
  // start at zero, also after a launch that was cut short.
  ACCEL_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s));
  const int grid = static_cast<int>((count + kStaticElemsPerCta - 1) / kStaticElemsPerCta);
  return grad_dtype == ACCEL_BF16
             ? launch_sumsq_static<bf16_t>(a, grid, partials, ticket, out_sumsq_dev, s)
             : launch_sumsq_static<float>(a, grid, partials, ticket, out_sumsq_dev, s);
}
