// Shared by the one-kernel forms of the multi-GPU PS step (ps_shard_fused.cu: unicast peer
// stores; ps_shard_switch.cu: in-switch reduction + multicast): argument block, the
// non-blocking poll, the closing ticket, and the cooperative launch.
#pragma once

#include <cooperative_groups.h>
#include <stdlib.h>

#include "ps_shard.cuh"

namespace accel {

constexpr int kFusedThreads = 256;
constexpr int kMaxSlices = 64;
constexpr int kEmuRanks = 8;

template <int R>
struct FusedArgs {
  ShardArgs r[R];
  int ctas_per_rank;
  int n_slices;
  int roles;  // ACCEL_OVERLAP_ROLES: percent of the CTAs that only apply + pull (0 = every CTA does both)
  int n_reducers;  // ACCEL_OVERLAP_SWITCH: CTAs per rank that reduce + apply + pull; the rest sum
  int trace;       // development: stamp %globaltimer into flag words [128, 256)
  int l2hint;      // development: L2 evict-first gradient loads / evict-last stage stores (summers)
  uint32_t epoch_base;  // slice k is published as epoch_base + k + 1, the pull as epoch_base + n_slices
};

// Non-blocking: have all ranks published >= target in words [base, base+world)?
__device__ inline bool poll_all(uint32_t *flags, int base, int world, uint32_t target) {
  __shared__ int ready;
  if (threadIdx.x == 0) {
    int ok = 1;
    for (int s = 0; s < world; ++s)
      ok &= static_cast<int32_t>(ld_relaxed_sys(flags + base + s) - target) >= 0;
    if (ok) __threadfence_system();   // the relaxed observations become an acquire
    ready = ok;
  }
  __syncthreads();
  const bool r = ready != 0;
  __syncthreads();  // `ready` may be rewritten by the next poll
  return r;
}


__device__ __forceinline__ Scalars shard_scalars(const ShardArgs &a) {
  ApplyArgs sa;  // scalars only
  sa.sumsq = a.sumsq;
  sa.dev_powers = nullptr;
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
  return make_scalars(sa);
}

// End of the one-kernel step. EVERY CTA of the rank arrives here exactly once — also after
// one of its waits timed out — so the ticket always returns to zero. The CTA that takes the
// last ticket publishes "rank r pulled" to every peer (unless a wait of this rank failed: its
// replica writes are then incomplete and the peers must see a timeout, not a completion) and
// waits (bounded) for the same from every peer: kernel completion = this rank's replica is
// complete and its stage may be overwritten.
__device__ inline void close_step(const ShardArgs &a, uint32_t *flags, int nb, uint32_t pull_epoch) {
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int done = atomicAdd(flags + kFlagTicketApply, 1u);
    last = done == static_cast<unsigned int>(nb) - 1;
    if (last) {
      flags[kFlagTicketApply] = 0;
      __threadfence_system();
      if (ld_relaxed_sys(flags + kFlagStatus) != 0) {
        last = 0;
      } else {
        for (int p = 0; p < a.world; ++p) st_relaxed_sys(a.flags[p] + kFlagPull + a.rank, pull_epoch);
      }
    }
  }
  __syncthreads();
  if (last) wait_all(flags, kFlagPull, a.world, pull_epoch);
}

// Grid of a one-kernel step: CTAs per rank such that the whole grid is co-resident, and no
// more CTAs than there is work for (small shards: fewer tickets and fences per step).
template <typename Args>
static int size_resident(void (*kern)(const Args), int n_ranks, const TaskState &ts,
                         int ctas_per_sm_cap, uint64_t work_items, int *per_rank_out,
                         int threads = kFusedThreads) {
  int per_sm = 0;
  ACCEL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, 0));
  if (per_sm < 1) return fail(ACCEL_ERR_CUDA, "one-kernel step does not fit on an SM");
  if (per_sm > ctas_per_sm_cap) per_sm = ctas_per_sm_cap;
  const int resident = per_sm * ts.sm_count;
  int per_rank = resident / n_ranks;
  // items per thread below which the grid shrinks (development knob ACCEL_STEP_ITEMS, default 2)
  static const int items_per_thread = [] {
    const char *e = getenv("ACCEL_STEP_ITEMS");
    const int v = e ? atoi(e) : 0;
    return v >= 1 && v <= 64 ? v : 2;
  }();
  const uint64_t per_cta = static_cast<uint64_t>(items_per_thread) * kFusedThreads;
  const uint64_t want = (work_items + per_cta - 1) / per_cta;
  if (static_cast<uint64_t>(per_rank) > want) per_rank = static_cast<int>(want < 2 ? 2 : want);
  if (per_rank < 1 || per_rank * n_ranks > resident)
    return fail(ACCEL_ERR_INVALID, "%d ranks do not fit in %d resident CTAs", n_ranks, resident);
  *per_rank_out = per_rank;
  return ACCEL_OK;
}

// Launch with all CTAs co-resident: cooperative by default; with coop = 0 and a grid of at most
// one CTA per SM a plain launch (lower latency; co-resident as long as nothing else occupies
// the GPU, and every wait is bounded).
template <typename Args>
static int launch_sized(void (*kern)(const Args), Args &fa, int n_ranks, const TaskState &ts,
                        const Tuning &tune, cudaStream_t s, int threads = kFusedThreads) {
  const int grid = fa.ctas_per_rank * n_ranks;
  if (!tune.coop && grid <= ts.sm_count) {
    kern<<<grid, threads, 0, s>>>(fa);
    ACCEL_CUDA(cudaGetLastError());
    return ACCEL_OK;
  }
  void *params[] = {&fa};
  ACCEL_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void *>(kern), dim3(grid),
                                         dim3(threads), params, 0, s));
  return ACCEL_OK;
}

template <typename Args>
static int launch_resident(void (*kern)(const Args), Args &fa, int n_ranks, const TaskState &ts,
                           int ctas_per_sm_cap, uint64_t work_items, const Tuning &tune,
                           cudaStream_t s) {
  int rc = size_resident<Args>(kern, n_ranks, ts, ctas_per_sm_cap, work_items, &fa.ctas_per_rank);
  if (rc) return rc;
  return launch_sized<Args>(kern, fa, n_ranks, ts, tune, s);
}

int fill_rank(const accel_peers *peers, int optimizer, float *slot0, float *slot1,
              const void *const *grads, int n_workers, int grad_dtype, int n_contrib,
              const accel_hparams *hp, const float *sumsq, ShardArgs *a);
int clamp_slices(int n_slices, uint64_t shard_elems);
// ps_shard_switch.cu
int launch_switch(FusedArgs<1> &fa, int optimizer, int grad_dtype, const TaskState &ts,
                  const Tuning &tune, cudaStream_t s);
int launch_switch_group(FusedArgs<kEmuRanks> &fa, int n_ranks, int optimizer, int grad_dtype,
                        const TaskState &ts, const Tuning &tune, cudaStream_t s);

}  // namespace accel
