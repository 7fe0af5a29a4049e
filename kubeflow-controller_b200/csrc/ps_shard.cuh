// Shared declarations of the multi-GPU PS step: argument block, flag layout, and the
// cross-GPU signalling helpers. Included by ps_shard.cu (three-phase kernels) and
// ps_shard_fused.cu (the one-kernel step).
#pragma once

#include <string.h>

#include "ps_math.cuh"

namespace accel {

constexpr int kFlagPush = 0;     // [0,16): push of rank s has landed (epoch)
constexpr int kFlagPull = 16;    // [16,32): apply_pull of rank s has landed
constexpr int kFlagTicketPush = 32;
constexpr int kFlagTicketApply = 33;
constexpr int kFlagSliceTicket = 64;  // [64,128): per-slice tickets of the one-kernel step
constexpr int kFlagStatus = 48;  // 0 ok, else error code | (waited-for rank << 8)
constexpr int kPushGranule = 256;  // items (8 elements each) sent to one owner before the next
constexpr unsigned long long kSpinTimeoutNs = 2000000000ull;  // 2 s

struct ShardArgs {
  float *stage[ACCEL_MAX_PS];
  float *var[ACCEL_MAX_PS];
  uint32_t *flags[ACCEL_MAX_PS];
  const void *grad[ACCEL_MAX_WORKERS];
  float *slot0;
  float *slot1;
  const float *sumsq;
  float *var_mc;   // multicast mapping of all ranks' var replicas, or nullptr
  float *stage_mc; // multicast mapping of all ranks' stage buffers (in-switch step), or nullptr
  int stage_bf16;  // stage slots hold raw bf16 (one bf16 gradient per rank); else fp32 partials
  int in_place;    // in-switch step: the single local gradient already IS this rank's stage buffer
  uint16_t *var_bf16[ACCEL_MAX_PS];  // bf16 replicas (pull target) or all nullptr
  uint64_t shard;  // elements per shard, multiple of 8
  uint64_t sub_begin, sub_count;  // the slice of EVERY shard this launch covers (multiples of 8)
  int rank, world, n_workers;
  uint32_t epoch;
  // optimizer scalars (see ApplyArgs)
  int use_nesterov;
  float lr, momentum, beta1, one_minus_beta1, one_minus_beta2, alpha, epsilon, mean_div, clip_norm;
};

__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t *p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Thread 0 of the CTA waits until flags[base + s] has reached `epoch` for every
// rank s. Returns false (for the whole CTA) on timeout or if another CTA
// already recorded a failure.
__device__ inline bool wait_all(uint32_t *flags, int base, int world, uint32_t epoch) {
  __shared__ int ok;
  if (threadIdx.x == 0) {
    ok = 1;
    const unsigned long long t0 = globaltimer_ns();
    for (int s = 0; s < world && ok; ++s) {
      // epochs only grow; signed distance handles the 32-bit wrap. Relaxed polls, then ONE
      // system-scope fence below turns the observations into an acquire.
      while (static_cast<int32_t>(ld_relaxed_sys(flags + base + s) - epoch) < 0) {
        if (globaltimer_ns() - t0 > kSpinTimeoutNs || ld_relaxed_sys(flags + kFlagStatus) != 0) {
          atomicCAS(flags + kFlagStatus, 0u, 1u | (static_cast<uint32_t>(s) << 8) |
                                                 (static_cast<uint32_t>(base) << 16));
          ok = 0;
          break;
        }
        __nanosleep(100);
      }
    }
    __threadfence_system();
  }
  __syncthreads();
  return ok != 0;
}

// After the CTA's stores: the last CTA of the grid publishes `epoch` into word
// [word_base + rank] of every rank's flags.
__device__ inline void signal_all(const ShardArgs &a, int ticket_word, int word_base) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();  // this CTA's peer stores are ordered before the ticket
    uint32_t *local = a.flags[a.rank];
    const unsigned int done = atomicAdd(local + ticket_word, 1u);
    if (done == gridDim.x - 1) {
      local[ticket_word] = 0;  // ready for the next launch
      __threadfence_system();  // all CTAs' tickets (and their stores) happen-before the flags
      // one fence, then relaxed stores: a release per peer would pay the fence `world` times
      for (int p = 0; p < a.world; ++p) st_relaxed_sys(a.flags[p] + word_base + a.rank, a.epoch);
    }
  }
}

__device__ __forceinline__ void st_256(float *p, const float (&o)[8]) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(o[0]), "f"(o[1]),
               "f"(o[2]), "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7])
               : "memory");
}
__device__ __forceinline__ void st_multimem_128(float *mc, const float *o) {
  asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(o[0]),
               "f"(o[1]), "f"(o[2]), "f"(o[3])
               : "memory");
}
// fp32 -> bf16, round to nearest even (finite values; NaN keeps a quiet-NaN pattern)
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void st_128u(uint16_t *p, const uint32_t (&o)[4]) {
  asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(o[0]), "r"(o[1]), "r"(o[2]),
               "r"(o[3])
               : "memory");
}
// The pull of one item (8 updated values) of rank `a.rank`'s shard.
__device__ __forceinline__ void pull_item(const ShardArgs &a, uint64_t var_elem, const float (&v)[8]) {
  if (a.var_bf16[0] != nullptr) {
    st_state_256(a.var[a.rank] + var_elem, v);      // fp32 master stays with the owner
    uint32_t h[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
    for (int q = 0; q < a.world; ++q) {
      int p = a.rank + q;
      if (p >= a.world) p -= a.world;
      st_128u(a.var_bf16[p] + var_elem, h);
    }
  } else if (a.var_mc != nullptr) {
    // through the switch: one multicast store lands in every rank's replica
    st_multimem_128(a.var_mc + var_elem, v);
    st_multimem_128(a.var_mc + var_elem + 4, v + 4);
  } else {
    // unicast peer stores: own replica first, then the right-hand neighbours
    for (int q = 0; q < a.world; ++q) {
      int p = a.rank + q;
      if (p >= a.world) p -= a.world;
      st_256(a.var[p] + var_elem, v);
    }
  }
}
__device__ __forceinline__ void ld_256_plain(const float *p, float (&o)[8]);
// 8 raw bf16 values (one 128-bit load) written by a peer during this step: coherent path
__device__ __forceinline__ void ld_128u_plain(const uint16_t *p, uint32_t (&o)[4]) {
  asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3])
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void widen_bf16x8(const uint32_t (&u)[4], float (&o)[8]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    o[2 * k] = bf16_lo(u[k]);
    o[2 * k + 1] = bf16_hi(u[k]);
  }
}
// One item (8 elements) of stage slot `slot` of shard-local item `it`, as fp32.
__device__ __forceinline__ void ld_stage_item(const ShardArgs &a, const float *stage, int slot,
                                              uint64_t shard_items, uint64_t it, float (&o)[8]) {
  const uint64_t idx = (static_cast<uint64_t>(slot) * shard_items + it) * 8;
  if (a.stage_bf16) {
    uint32_t u[4];
    ld_128u_plain(reinterpret_cast<const uint16_t *>(stage) + idx, u);
    widen_bf16x8(u, o);
  } else {
    ld_256_plain(stage + idx, o);
  }
}
__device__ __forceinline__ void ld_256_plain(const float *p, float (&o)[8]) {
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]), "=f"(o[4]), "=f"(o[5]),
                 "=f"(o[6]), "=f"(o[7])
               : "l"(p)
               : "memory");
}


// host side, defined in ps_shard.cu
int fill_peers(const accel_peers *peers, uint64_t sub_begin, uint64_t sub_count, ShardArgs *a);
int fill_optimizer(int optimizer, float *slot0, float *slot1, int n_contrib,
                   const accel_hparams *hp, const float *grad_sumsq_dev, ShardArgs *a);
int fill_grads(const void *const *grads, int n_workers, int grad_dtype, ShardArgs *a);

}  // namespace accel
