// Multi-GPU PS step over NVLink peer memory (SURVEY §8 e): PS task r = GPU r =
// rank r owns shard r of the flat arena. One process per GPU; peers' buffers are
// mapped with CUDA IPC. Two fused kernels per step replace the reference's gRPC
// push/pull between pods (pkg/tensorflow/distributed.go:26-29, port 2222):
//
//   push        : local N-way gradient sum fused with the reduce-scatter — the
//                 partial for shard c is stored straight into rank c's stage
//                 slot [rank] (peer store over NVLink), so the transfer rides
//                 under the HBM-bound read of the local gradients;
//   apply_pull  : shard owner sums the `world` stage slots in rank order,
//                 applies the optimizer in registers and stores the updated
//                 values into every rank's variable replica (the all-gather).
//
// Cross-GPU ordering uses epoch flags in each rank's own memory, written by the
// peers with release stores at system scope after a last-CTA ticket, and polled
// with acquire loads. Every poll is bounded (kSpinTimeoutNs): on timeout the
// kernel records an error in the status word and returns without touching data.
#include "ps_math.cuh"

namespace accel {

constexpr int kFlagPush = 0;     // [0,16): push of rank s has landed (epoch)
constexpr int kFlagPull = 16;    // [16,32): apply_pull of rank s has landed
constexpr int kFlagTicketPush = 32;
constexpr int kFlagTicketApply = 33;
constexpr int kFlagStatus = 48;  // 0 ok, else error code | (waited-for rank << 8)
constexpr unsigned long long kSpinTimeoutNs = 2000000000ull;  // 2 s

struct ShardArgs {
  float *stage[ACCEL_MAX_PS];
  float *var[ACCEL_MAX_PS];
  uint32_t *flags[ACCEL_MAX_PS];
  const void *grad[ACCEL_MAX_WORKERS];
  float *slot0;
  float *slot1;
  const float *sumsq;
  uint64_t shard;  // elements per shard, multiple of 8
  int rank, world, n_workers;
  uint32_t epoch;
  // optimizer scalars (see ApplyArgs)
  int use_nesterov;
  float lr, momentum, beta1, one_minus_beta1, one_minus_beta2, alpha, epsilon, mean_div, clip_norm;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Thread 0 of the CTA waits until flags[base + s] has reached `epoch` for every
// rank s. Returns false (for the whole CTA) on timeout or if another CTA
// already recorded a failure.
__device__ bool wait_all(uint32_t *flags, int base, int world, uint32_t epoch) {
  __shared__ int ok;
  if (threadIdx.x == 0) {
    ok = 1;
    const unsigned long long t0 = globaltimer_ns();
    for (int s = 0; s < world && ok; ++s) {
      // epochs only grow; signed distance handles the 32-bit wrap
      while (static_cast<int32_t>(ld_acquire_sys(flags + base + s) - epoch) < 0) {
        if (globaltimer_ns() - t0 > kSpinTimeoutNs || ld_acquire_sys(flags + kFlagStatus) != 0) {
          atomicCAS(flags + kFlagStatus, 0u, 1u | (static_cast<uint32_t>(s) << 8) |
                                                 (static_cast<uint32_t>(base) << 16));
          ok = 0;
          break;
        }
        __nanosleep(200);
      }
    }
  }
  __syncthreads();
  return ok != 0;
}

// After the CTA's stores: the last CTA of the grid publishes `epoch` into word
// [word_base + rank] of every rank's flags.
__device__ void signal_all(const ShardArgs &a, int ticket_word, int word_base) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();  // this CTA's peer stores are ordered before the ticket
    uint32_t *local = a.flags[a.rank];
    const unsigned int done = atomicAdd(local + ticket_word, 1u);
    if (done == gridDim.x - 1) {
      local[ticket_word] = 0;  // ready for the next launch
      __threadfence_system();  // all CTAs' tickets (and their stores) happen-before the flags
      for (int p = 0; p < a.world; ++p) st_release_sys(a.flags[p] + word_base + a.rank, a.epoch);
    }
  }
}

__device__ __forceinline__ void st_256(float *p, const float (&o)[8]) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(o[0]), "f"(o[1]),
               "f"(o[2]), "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7])
               : "memory");
}
__device__ __forceinline__ void ld_256_plain(const float *p, float (&o)[8]) {
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]), "=f"(o[4]), "=f"(o[5]),
                 "=f"(o[6]), "=f"(o[7])
               : "l"(p)
               : "memory");
}

// ---- push: local N-way sum fused with the reduce-scatter -----------------------
// Items of 8 elements. Logical item j maps to arena element
// ((rank+1)*shard + 8j) mod (world*shard): every rank starts with its right-hand
// neighbour's shard, so at any moment the `world` ranks store to `world` different
// destinations and no GPU's NVLink ingress is the hot spot.
template <typename G, int NW, int U>
__global__ void __launch_bounds__(512) ps_push(const __grid_constant__ ShardArgs a) {
  const uint64_t shard_items = a.shard / 8;
  const uint64_t n_items = shard_items * a.world;
  const uint64_t rot = shard_items * ((a.rank + 1) % a.world);
  const uint64_t tile = static_cast<uint64_t>(blockDim.x) * U;
  const int n = NW > 0 ? NW : a.n_workers;

  for (uint64_t base = static_cast<uint64_t>(blockIdx.x) * tile; base < n_items;
       base += static_cast<uint64_t>(gridDim.x) * tile) {
    uint64_t item[U];
    bool ok[U];
    float acc[U][8];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const uint64_t l = base + static_cast<uint64_t>(j) * blockDim.x + threadIdx.x;
      ok[j] = l < n_items;
      uint64_t p = l + rot;
      item[j] = p >= n_items ? p - n_items : p;
    }
    if constexpr (NW > 0) {
      float g[NW][U][8];
#pragma unroll
      for (int w = 0; w < NW; ++w)
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j]) load_grad_item<G, true>(a.grad[w], item[j], g[w][j]);
#pragma unroll
      for (int j = 0; j < U; ++j)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          acc[j][e] = g[0][j][e];
#pragma unroll
          for (int w = 1; w < NW; ++w) acc[j][e] = __fadd_rn(acc[j][e], g[w][j][e]);
        }
    } else {
#pragma unroll
      for (int j = 0; j < U; ++j)
        if (ok[j]) load_grad_item<G, true>(a.grad[0], item[j], acc[j]);
      for (int w = 1; w < n; ++w) {
        float g[U][8];
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j]) load_grad_item<G, true>(a.grad[w], item[j], g[j]);
#pragma unroll
        for (int j = 0; j < U; ++j)
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[j][e] = __fadd_rn(acc[j][e], g[j][e]);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      if (!ok[j]) continue;
      const uint64_t c = item[j] / shard_items;           // destination rank = shard owner
      const uint64_t in_shard = item[j] - c * shard_items;
      float *dst = a.stage[c] + (static_cast<uint64_t>(a.rank) * shard_items + in_shard) * 8;
      st_256(dst, acc[j]);
    }
  }
  signal_all(a, kFlagTicketPush, kFlagPush);
}

// ---- apply + pull: owner applies its shard and broadcasts the new values ---------
template <int OPT, int U>
__global__ void __launch_bounds__(512) ps_apply_pull(const __grid_constant__ ShardArgs a) {
  constexpr int SLOTS = OptTraits<OPT>::kSlots;
  uint32_t *flags = a.flags[a.rank];
  if (!wait_all(flags, kFlagPush, a.world, a.epoch)) return;

  ApplyArgs sa;  // scalars only; make_scalars reads nothing else
  sa.sumsq = a.sumsq;
  sa.use_nesterov = a.use_nesterov;
  sa.lr = a.lr;
  sa.momentum = a.momentum;
  sa.beta1 = a.beta1;
  sa.one_minus_beta1 = a.one_minus_beta1;
  sa.one_minus_beta2 = a.one_minus_beta2;
  sa.alpha = a.alpha;
  sa.epsilon = a.epsilon;
  sa.mean_div = a.mean_div;
  sa.clip_norm = a.clip_norm;
  const Scalars sc = make_scalars(sa);

  const uint64_t shard_items = a.shard / 8;
  const float *stage = a.stage[a.rank];
  const uint64_t var_off = static_cast<uint64_t>(a.rank) * a.shard;
  const uint64_t tile = static_cast<uint64_t>(blockDim.x) * U;

  for (uint64_t base = static_cast<uint64_t>(blockIdx.x) * tile; base < shard_items;
       base += static_cast<uint64_t>(gridDim.x) * tile) {
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const uint64_t it = base + static_cast<uint64_t>(j) * blockDim.x + threadIdx.x;
      if (it >= shard_items) continue;
      float acc[8], v[8], s0[8], s1[8];
      ld_state_256(a.var[a.rank] + var_off + it * 8, v);
      if constexpr (SLOTS >= 1) ld_state_256(a.slot0 + it * 8, s0);
      if constexpr (SLOTS >= 2) ld_state_256(a.slot1 + it * 8, s1);
      // stage slots in rank order, four loads in flight at a time
      for (int s0i = 0; s0i < a.world; s0i += 4) {
        float g[4][8];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (s0i + k < a.world)
            ld_256_plain(stage + (static_cast<uint64_t>(s0i + k) * shard_items + it) * 8, g[k]);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (s0i + k < a.world)
#pragma unroll
            for (int e = 0; e < 8; ++e)
              acc[e] = (s0i + k == 0) ? g[k][e] : __fadd_rn(acc[e], g[k][e]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) update<OPT>(v[e], s0[e], s1[e], finish_grad(acc[e], sc), sc);
      if constexpr (SLOTS >= 1) st_state_256(a.slot0 + it * 8, s0);
      if constexpr (SLOTS >= 2) st_state_256(a.slot1 + it * 8, s1);
      // the pull: own replica first, then the peers starting with the right-hand neighbour
      for (int k = 0; k < a.world; ++k) {
        int p = a.rank + k;
        if (p >= a.world) p -= a.world;
        st_256(a.var[p] + var_off + it * 8, v);
      }
    }
  }
  signal_all(a, kFlagTicketApply, kFlagPull);
}

__global__ void ps_wait_pull(const __grid_constant__ ShardArgs a) {
  wait_all(a.flags[a.rank], kFlagPull, a.world, a.epoch);
}

template <typename G>
static int launch_push(const ShardArgs &a, int grid, cudaStream_t s) {
  const int threads = 256;
#define ACCEL_PUSH(NW, U) ps_push<G, NW, U><<<grid, threads, 0, s>>>(a)
  switch (a.n_workers) {
    case 1: ACCEL_PUSH(1, 2); break;
    case 2: ACCEL_PUSH(2, 2); break;
    case 4: ACCEL_PUSH(4, 2); break;
    case 8: ACCEL_PUSH(8, 1); break;
    default: ACCEL_PUSH(0, 2); break;
  }
#undef ACCEL_PUSH
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

static int fill_peers(const accel_peers *peers, ShardArgs *a) {
  if (!peers) return fail(ACCEL_ERR_INVALID, "peers is NULL");
  if (peers->world < 1 || peers->world > ACCEL_MAX_PS)
    return fail(ACCEL_ERR_INVALID, "world %d out of range [1,%d]", peers->world, ACCEL_MAX_PS);
  if (peers->rank < 0 || peers->rank >= peers->world)
    return fail(ACCEL_ERR_INVALID, "rank %d out of range [0,%d)", peers->rank, peers->world);
  if (peers->shard_elems == 0 || peers->shard_elems % 8)
    return fail(ACCEL_ERR_INVALID, "shard_elems %llu must be a positive multiple of 8",
                static_cast<unsigned long long>(peers->shard_elems));
  if (!peers->stage || !peers->var || !peers->flags)
    return fail(ACCEL_ERR_INVALID, "peers->stage/var/flags is NULL");
  memset(a, 0, sizeof(*a));
  for (int r = 0; r < peers->world; ++r) {
    if (!peers->stage[r] || !peers->var[r] || !peers->flags[r])
      return fail(ACCEL_ERR_INVALID, "peer %d has a NULL buffer", r);
    if ((reinterpret_cast<uintptr_t>(peers->stage[r]) | reinterpret_cast<uintptr_t>(peers->var[r])) & 31)
      return fail(ACCEL_ERR_ALIGN, "peer %d stage/var must be 32-byte aligned", r);
    a->stage[r] = peers->stage[r];
    a->var[r] = peers->var[r];
    a->flags[r] = peers->flags[r];
  }
  a->rank = peers->rank;
  a->world = peers->world;
  a->shard = peers->shard_elems;
  return ACCEL_OK;
}

}  // namespace accel

using namespace accel;

extern "C" {

int accel_dev_alloc(size_t bytes, void **out) {
  if (!out || !bytes) return fail(ACCEL_ERR_INVALID, "out is NULL or bytes is 0");
  ACCEL_CUDA(cudaMalloc(out, bytes));
  ACCEL_CUDA(cudaMemset(*out, 0, bytes));
  return ACCEL_OK;
}

int accel_dev_free(void *ptr) {
  if (ptr) ACCEL_CUDA(cudaFree(ptr));
  return ACCEL_OK;
}

int accel_shard_push(accel_ctx *ctx, const accel_peers *peers, const void *const *grads,
                     int n_workers, int grad_dtype, accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  if (!grads) return fail(ACCEL_ERR_INVALID, "grads is NULL");
  if (n_workers < 1 || n_workers > ACCEL_MAX_WORKERS)
    return fail(ACCEL_ERR_INVALID, "n_workers %d out of range [1,%d]", n_workers,
                ACCEL_MAX_WORKERS);
  if (grad_dtype != ACCEL_F32 && grad_dtype != ACCEL_BF16)
    return fail(ACCEL_ERR_INVALID, "unknown grad_dtype %d", grad_dtype);
  for (int w = 0; w < n_workers; ++w) {
    if (!grads[w]) return fail(ACCEL_ERR_INVALID, "grads[%d] is NULL", w);
    if (reinterpret_cast<uintptr_t>(grads[w]) & 31)
      return fail(ACCEL_ERR_ALIGN, "grads[%d] must be 32-byte aligned", w);
    a.grad[w] = grads[w];
  }
  a.n_workers = n_workers;
  a.epoch = ++ts->shard_epoch;  // one epoch per step: push, apply_pull and wait_pull share it
  const uint64_t items = a.shard / 8 * a.world;
  int grid = ts->sm_count * 2;
  const uint64_t want = (items + 255) / 256;
  if (static_cast<uint64_t>(grid) > want) grid = static_cast<int>(want);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return grad_dtype == ACCEL_BF16 ? launch_push<bf16_t>(a, grid, s) : launch_push<float>(a, grid, s);
}

int accel_shard_apply_pull(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                           float *slot1, int n_contrib, const accel_hparams *hp,
                           const float *grad_sumsq_dev, accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  if (!hp) return fail(ACCEL_ERR_INVALID, "hparams is NULL");
  if (optimizer < ACCEL_OPT_SGD || optimizer > ACCEL_OPT_ADAM)
    return fail(ACCEL_ERR_INVALID, "unknown optimizer %d", optimizer);
  if (optimizer >= ACCEL_OPT_MOMENTUM && !slot0)
    return fail(ACCEL_ERR_INVALID, "optimizer slot 0 (accum / m) is NULL");
  if (optimizer == ACCEL_OPT_ADAM && !slot1)
    return fail(ACCEL_ERR_INVALID, "optimizer slot 1 (v) is NULL");
  if ((reinterpret_cast<uintptr_t>(slot0) | reinterpret_cast<uintptr_t>(slot1)) & 31)
    return fail(ACCEL_ERR_ALIGN, "slots must be 32-byte aligned");
  if (hp->clip_norm > 0.0f && !grad_sumsq_dev)
    return fail(ACCEL_ERR_INVALID,
                "clip_norm > 0 needs grad_sumsq_dev: the global norm spans all shards");
  if (n_contrib < 1) return fail(ACCEL_ERR_INVALID, "n_contrib must be >= 1");
  if (ts->shard_epoch == 0)
    return fail(ACCEL_ERR_STATE, "accel_shard_apply_pull before any accel_shard_push");
  a.slot0 = slot0;
  a.slot1 = slot1;
  a.sumsq = grad_sumsq_dev;
  a.epoch = ts->shard_epoch;
  a.use_nesterov = hp->use_nesterov ? 1 : 0;
  a.lr = hp->lr;
  a.momentum = hp->momentum;
  a.beta1 = hp->beta1;
  a.one_minus_beta1 = 1.0f - hp->beta1;
  a.one_minus_beta2 = 1.0f - hp->beta2;
  a.epsilon = hp->epsilon;
  if (optimizer == ACCEL_OPT_ADAM) {
    const float num = hp->lr * sqrtf(1.0f - hp->beta2_power);
    a.alpha = num / (1.0f - hp->beta1_power);
  }
  a.mean_div = hp->aggregate == ACCEL_AGG_MEAN ? static_cast<float>(n_contrib) : 0.0f;
  a.clip_norm = hp->clip_norm > 0.0f ? hp->clip_norm : 0.0f;
  const uint64_t items = a.shard / 8;
  int grid = ts->sm_count * 4;
  const uint64_t want = (items + 255) / 256;
  if (static_cast<uint64_t>(grid) > want) grid = static_cast<int>(want);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (optimizer) {
    case ACCEL_OPT_SGD: ps_apply_pull<OPT_SGD, 1><<<grid, 256, 0, s>>>(a); break;
    case ACCEL_OPT_MOMENTUM: ps_apply_pull<OPT_MOMENTUM, 1><<<grid, 256, 0, s>>>(a); break;
    default: ps_apply_pull<OPT_ADAM, 1><<<grid, 256, 0, s>>>(a); break;
  }
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

int accel_shard_wait_pull(accel_ctx *ctx, const accel_peers *peers, accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  if (ts->shard_epoch == 0)
    return fail(ACCEL_ERR_STATE, "accel_shard_wait_pull before any accel_shard_push");
  a.epoch = ts->shard_epoch;
  ps_wait_pull<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a);
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

int accel_shard_check(accel_ctx *ctx, const accel_peers *peers, accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  ACCEL_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  uint32_t status = 0;
  ACCEL_CUDA(cudaMemcpy(&status, a.flags[a.rank] + kFlagStatus, sizeof(status),
                        cudaMemcpyDeviceToHost));
  if (status != 0)
    return fail(ACCEL_ERR_STATE,
                "rank %d timed out after %.1f s waiting for the %s signal of rank %u (epoch %u); "
                "the step's results are invalid",
                a.rank, kSpinTimeoutNs * 1e-9, ((status >> 16) & 0xff) == kFlagPush ? "push" : "pull",
                (status >> 8) & 0xff, ts->shard_epoch);
  return ACCEL_OK;
}

}  // extern "C"
