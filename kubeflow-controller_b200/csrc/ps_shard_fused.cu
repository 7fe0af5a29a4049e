// The multi-GPU PS step as ONE kernel per GPU (SURVEY §8 e): local N-way gradient sum +
// reduce-scatter push, shard apply, all-gather pull, and the closing wait, overlapped
// slice by slice over NVLink peer memory.
//
// Every CTA walks the slices in order. For slice k it (1) sums its part of the local
// gradients and stores the partials into the owners' stage slots (peer stores are posted:
// the SM moves on while NVLink drains), (2) takes a per-slice ticket — the last CTA
// publishes "rank r pushed slice k" to every peer — and (3) without blocking, checks
// whether earlier slices have arrived from ALL ranks and, if so, applies the optimizer to
// its part of them and stores the new values into every replica. Only after its last push
// does a CTA block (bounded) for the slices still missing. The last CTA to finish
// publishes "rank r pulled" and waits (bounded) for the same from every peer, so kernel
// completion = this rank's replica is complete.
//
// Progress: a CTA only ever BLOCKS after all of its own pushes are issued, and pushes never
// wait, so every rank's push flags eventually arrive provided all CTAs of the grid are
// resident — the launch is cooperative (co-residency guaranteed or the launch fails).
//
// R > 1 emulates R ranks that live on ONE GPU inside a single cooperative launch (CTA group
// g plays rank g): the single-GPU test vehicle for the same code path.
#include "ps_shard_fused.cuh"

namespace accel {

// One item per thread and iteration: the form for 4, 8 or a run-time number of gradients
// (registers: 8 gradients x 8 values are already 64 in flight).
template <typename G, int NW>
__device__ __forceinline__ void push_slice_one(const ShardArgs &a, uint64_t sub_first,
                                           uint64_t sub_items, int bid, int nb) {
  const uint64_t shard_items = a.shard / 8;
  const uint64_t granules = (sub_items + kPushGranule - 1) / kPushGranule;   // per owner
  const uint64_t n_items = granules * kPushGranule * a.world;
  const int n = NW > 0 ? NW : a.n_workers;
  for (uint64_t l = static_cast<uint64_t>(bid) * blockDim.x + threadIdx.x; l < n_items;
       l += static_cast<uint64_t>(nb) * blockDim.x) {
    // granule g goes to owner (rank + 1 + g) mod world: all owners are written all the time
    const uint64_t g_idx = l / kPushGranule;
    const uint64_t in_shard = (g_idx / a.world) * kPushGranule + (l % kPushGranule);
    if (in_shard >= sub_items) continue;
    int c = a.rank + 1 + static_cast<int>(g_idx % a.world);
    if (c >= a.world) c -= a.world;
    const uint64_t within = sub_first + in_shard;
    const uint64_t item = static_cast<uint64_t>(c) * shard_items + within;
    if constexpr (sizeof(G) == 2 && NW == 1) {
      if (a.stage_bf16) {   // one bf16 gradient per rank: forward the raw item (2 B/elem)
        uint32_t raw[4];
        ldg_stream_128u(static_cast<const uint16_t *>(a.grad[0]) + item * 8, raw);
        st_128u(reinterpret_cast<uint16_t *>(a.stage[c]) +
                    (static_cast<uint64_t>(a.rank) * shard_items + within) * 8, raw);
        continue;
      }
    }
    float acc[8];
    if constexpr (NW > 0) {
      float g[NW][8];
#pragma unroll
      for (int w = 0; w < NW; ++w) load_grad_item<G, true>(a.grad[w], item, g[w]);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        acc[e] = g[0][e];
#pragma unroll
        for (int w = 1; w < NW; ++w) acc[e] = __fadd_rn(acc[e], g[w][e]);
      }
    } else {
      load_grad_item<G, true>(a.grad[0], item, acc);
      for (int w0 = 1; w0 < n; w0 += 4) {
        float g[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (w0 + q < n) load_grad_item<G, true>(a.grad[w0 + q], item, g[q]);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (w0 + q < n)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = __fadd_rn(acc[e], g[q][e]);
      }
    }
    st_256(a.stage[c] + (static_cast<uint64_t>(a.rank) * shard_items + within) * 8, acc);
  }
}

template <typename G, int NW>
__device__ __forceinline__ void push_slice(const ShardArgs &a, uint64_t sub_first,
                                           uint64_t sub_items, int bid, int nb) {
  // items in flight per thread: the 4- and 8-gradient forms already have that many loads
  // outstanding; with one or two gradient streams a thread takes several items at once so that
  // neither HBM latency nor the link's is exposed (matters most for small shards)
  constexpr int U = (NW == 1) ? 2 : 1;   // 4 spills under the 128-register cap of 2 CTAs/SM
  if constexpr (U == 1) {
    push_slice_one<G, NW>(a, sub_first, sub_items, bid, nb);
    return;
  }
  const uint64_t shard_items = a.shard / 8;
  const uint64_t granules = (sub_items + kPushGranule - 1) / kPushGranule;   // per owner
  const uint64_t n_items = granules * kPushGranule * a.world;
  const int n = NW > 0 ? NW : a.n_workers;
  const uint64_t stride = static_cast<uint64_t>(nb) * blockDim.x;
  for (uint64_t l0 = static_cast<uint64_t>(bid) * blockDim.x + threadIdx.x; l0 < n_items;
       l0 += stride * U) {
    uint64_t item[U], dst[U];
    int owner[U];
    bool ok[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const uint64_t l = l0 + stride * j;
      // granule g goes to owner (rank + 1 + g) mod world: all owners are written all the time
      const uint64_t g_idx = l / kPushGranule;
      const uint64_t in_shard = (g_idx / a.world) * kPushGranule + (l % kPushGranule);
      ok[j] = l < n_items && in_shard < sub_items;
      int c = a.rank + 1 + static_cast<int>(g_idx % a.world);
      if (c >= a.world) c -= a.world;
      owner[j] = ok[j] ? c : a.rank;
      const uint64_t within = sub_first + (ok[j] ? in_shard : 0);
      item[j] = static_cast<uint64_t>(owner[j]) * shard_items + within;
      dst[j] = static_cast<uint64_t>(a.rank) * shard_items + within;
    }
    if constexpr (sizeof(G) == 2 && NW == 1) {
      if (a.stage_bf16) {   // one bf16 gradient per rank: forward the raw item (2 B/elem)
        uint32_t raw[U][4];
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j]) ldg_stream_128u(static_cast<const uint16_t *>(a.grad[0]) + item[j] * 8, raw[j]);
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j]) st_128u(reinterpret_cast<uint16_t *>(a.stage[owner[j]]) + dst[j] * 8, raw[j]);
        continue;
      }
    }
    float acc[U][8];
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
      if (ok[0]) load_grad_item<G, true>(a.grad[0], item[0], acc[0]);
      for (int w0 = 1; w0 < n; w0 += 4) {
        float g[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (w0 + q < n && ok[0]) load_grad_item<G, true>(a.grad[w0 + q], item[0], g[q]);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (w0 + q < n)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[0][e] = __fadd_rn(acc[0][e], g[q][e]);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j)
      if (ok[j]) st_256(a.stage[owner[j]] + dst[j] * 8, acc[j]);
  }
}

template <int OPT>
__device__ __forceinline__ void apply_slice(const ShardArgs &a, const Scalars &sc,
                                            uint64_t sub_first, uint64_t sub_items, int bid,
                                            int nb) {
  constexpr int SLOTS = OptTraits<OPT>::kSlots;
  const uint64_t shard_items = a.shard / 8;
  const float *stage = a.stage[a.rank];
  const uint64_t var_off = static_cast<uint64_t>(a.rank) * a.shard;
  for (uint64_t i = static_cast<uint64_t>(bid) * blockDim.x + threadIdx.x; i < sub_items;
       i += static_cast<uint64_t>(nb) * blockDim.x) {
    const uint64_t it = sub_first + i;
    float acc[8], v[8], s0[8], s1[8];
    ld_state_256(a.var[a.rank] + var_off + it * 8, v);
    if constexpr (SLOTS >= 1) ld_state_256(a.slot0 + it * 8, s0);
    if constexpr (SLOTS >= 2) ld_state_256(a.slot1 + it * 8, s1);
    for (int s0i = 0; s0i < a.world; s0i += 4) {
      float g[4][8];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (s0i + q < a.world) ld_stage_item(a, stage, s0i + q, shard_items, it, g[q]);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (s0i + q < a.world)
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = (s0i + q == 0) ? g[q][e] : __fadd_rn(acc[e], g[q][e]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) update<OPT>(v[e], s0[e], s1[e], finish_grad(acc[e], sc), sc);
    if constexpr (SLOTS >= 1) st_state_256(a.slot0 + it * 8, s0);
    if constexpr (SLOTS >= 2) st_state_256(a.slot1 + it * 8, s1);
    pull_item(a, var_off + it * 8, v);
  }
}

template <typename G, int NW, int OPT, int R>
__global__ void __launch_bounds__(kFusedThreads, 2)
    ps_step_fused(const __grid_constant__ FusedArgs<R> fa) {
  const int vr = R == 1 ? 0 : blockIdx.x / fa.ctas_per_rank;
  const ShardArgs &a = fa.r[vr];
  const int nb = fa.ctas_per_rank;
  const int bid = R == 1 ? blockIdx.x : blockIdx.x - vr * nb;
  uint32_t *flags = a.flags[a.rank];

  const Scalars sc = shard_scalars(a);

  const uint64_t shard_items = a.shard / 8;
  const int S = fa.n_slices;
  const uint64_t per = (shard_items + S - 1) / S;
  auto slice_first = [&](int k) { return per * k < shard_items ? per * k : shard_items; };
  auto slice_items = [&](int k) {
    const uint64_t f = slice_first(k);
    return f + per <= shard_items ? per : shard_items - f;
  };

  if (fa.roles && nb > 1) {
    // ---- split roles: pushers never stop to apply, appliers trail them by one slice --------
    // applier share in percent (fa.roles): 50 = every second CTA (the original split); any other
    // value puts the first n_apply CTAs on apply + pull and the rest on push
    int n_apply, rb;
    bool pusher;
    if (fa.roles == 50) {
      n_apply = nb / 2;
      rb = bid >> 1;
      pusher = (bid & 1) == 0;
    } else {
      n_apply = nb * fa.roles / 100;
      if (n_apply < 1) n_apply = 1;
      if (n_apply > nb - 1) n_apply = nb - 1;
      pusher = bid >= n_apply;
      rb = pusher ? bid - n_apply : bid;
    }
    const int n_push = nb - n_apply;
    if (pusher) {
      // Pusher. The per-slice ticket is taken per WARP (no CTA barrier): a warp that has
      // finished its part of slice k fences its own stores and moves on to slice k+1 while the
      // other warps keep the memory system busy.
      const unsigned int total_warps = static_cast<unsigned int>(n_push) * (blockDim.x >> 5);
      for (int k = 0; k < S; ++k) {
        push_slice<G, NW>(a, slice_first(k), slice_items(k), rb, n_push);
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
          __threadfence_system();
          const unsigned int done = atomicAdd(flags + kFlagSliceTicket + k, 1u);
          if (done == total_warps - 1) {
            flags[kFlagSliceTicket + k] = 0;
            __threadfence_system();
            for (int p = 0; p < a.world; ++p)
              st_relaxed_sys(a.flags[p] + kFlagPush + a.rank, fa.epoch_base + k + 1);
          }
        }
      }
    } else {
      // Applier: blocks (bounded) only on OTHER CTAs' pushes, which never block.
      for (int k = 0; k < S; ++k) {
        if (!wait_all(flags, kFlagPush, a.world, fa.epoch_base + k + 1)) break;   // timed out
        apply_slice<OPT>(a, sc, slice_first(k), slice_items(k), rb, n_apply);
      }
    }
  } else {
    // ---- every CTA does both: push slice k, then apply whatever has already arrived --------
    int next_apply = 0;
    for (int k = 0; k < S; ++k) {
      push_slice<G, NW>(a, slice_first(k), slice_items(k), bid, nb);
      // per-slice ticket: the last CTA of this rank publishes slice k to every rank
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned int done = atomicAdd(flags + kFlagSliceTicket + k, 1u);
        if (done == static_cast<unsigned int>(nb) - 1) {
          flags[kFlagSliceTicket + k] = 0;
          __threadfence_system();
          for (int p = 0; p < a.world; ++p)
            st_relaxed_sys(a.flags[p] + kFlagPush + a.rank, fa.epoch_base + k + 1);
        }
      }
      // whatever has already arrived from everybody gets applied now, without waiting
      while (next_apply < k &&
             poll_all(flags, kFlagPush, a.world, fa.epoch_base + next_apply + 1)) {
        apply_slice<OPT>(a, sc, slice_first(next_apply), slice_items(next_apply), bid, nb);
        ++next_apply;
      }
    }
    // all pushes of this CTA are issued: now it may block for what is still missing
    for (; next_apply < S; ++next_apply) {
      if (!wait_all(flags, kFlagPush, a.world, fa.epoch_base + next_apply + 1)) break;   // timed out
      apply_slice<OPT>(a, sc, slice_first(next_apply), slice_items(next_apply), bid, nb);
    }
  }
  // pull signal, then the closing wait by the CTA that took the last ticket
  close_step(a, flags, nb, fa.epoch_base + S);
}

template <typename G, int NW, int OPT, int R>
static int launch_fused_one(const FusedArgs<R> &fa_in, int n_ranks, const TaskState &ts,
                            const Tuning &tune, cudaStream_t s) {
  FusedArgs<R> fa = fa_in;
  const uint64_t items = fa.r[0].shard / 8 * fa.r[0].world;
  // coop = 0: one CTA per SM at most, so the plain launch is possible
  return launch_resident<FusedArgs<R>>(ps_step_fused<G, NW, OPT, R>, fa, n_ranks, ts,
                                       tune.coop ? 2 : 1, items, tune, s);
}

template <typename G, int OPT, int R>
static int launch_fused_nw(const FusedArgs<R> &fa, int n_ranks, const TaskState &ts,
                           const Tuning &cap, cudaStream_t s) {
  switch (fa.r[0].n_workers) {
    case 1: return launch_fused_one<G, 1, OPT, R>(fa, n_ranks, ts, cap, s);
    case 2: return launch_fused_one<G, 2, OPT, R>(fa, n_ranks, ts, cap, s);
    case 4: return launch_fused_one<G, 4, OPT, R>(fa, n_ranks, ts, cap, s);
    case 8: return launch_fused_one<G, 8, OPT, R>(fa, n_ranks, ts, cap, s);
    default: return launch_fused_one<G, 0, OPT, R>(fa, n_ranks, ts, cap, s);
  }
}

template <int R>
static int launch_fused(const FusedArgs<R> &fa, int n_ranks, int optimizer, int grad_dtype,
                        const TaskState &ts, const Tuning &cap, cudaStream_t s) {
  const bool bf = grad_dtype == ACCEL_BF16;
  switch (optimizer) {
    case ACCEL_OPT_SGD:
      return bf ? launch_fused_nw<bf16_t, OPT_SGD, R>(fa, n_ranks, ts, cap, s)
                : launch_fused_nw<float, OPT_SGD, R>(fa, n_ranks, ts, cap, s);
    case ACCEL_OPT_MOMENTUM:
      return bf ? launch_fused_nw<bf16_t, OPT_MOMENTUM, R>(fa, n_ranks, ts, cap, s)
                : launch_fused_nw<float, OPT_MOMENTUM, R>(fa, n_ranks, ts, cap, s);
    default:
      return bf ? launch_fused_nw<bf16_t, OPT_ADAM, R>(fa, n_ranks, ts, cap, s)
                : launch_fused_nw<float, OPT_ADAM, R>(fa, n_ranks, ts, cap, s);
  }
}

int fill_rank(const accel_peers *peers, int optimizer, float *slot0, float *slot1,
                     const void *const *grads, int n_workers, int grad_dtype, int n_contrib,
                     const accel_hparams *hp, const float *sumsq, ShardArgs *a) {
  int rc = fill_peers(peers, 0, 0, a);
  if (rc) return rc;
  rc = fill_grads(grads, n_workers, grad_dtype, a);
  if (rc) return rc;
  return fill_optimizer(optimizer, slot0, slot1, n_contrib, hp, sumsq, a);
}

int clamp_slices(int n_slices, uint64_t shard_elems) {
  if (n_slices < 1) n_slices = 1;
  if (n_slices > kMaxSlices) n_slices = kMaxSlices;
  const uint64_t items = shard_elems / 8;
  if (static_cast<uint64_t>(n_slices) > items) n_slices = static_cast<int>(items);
  return n_slices;
}

}  // namespace accel

using namespace accel;

static bool overlap_ok(int overlap) {
  return overlap == ACCEL_OVERLAP_INLINE || overlap == ACCEL_OVERLAP_ROLES ||
         overlap == ACCEL_OVERLAP_SWITCH;
}

// SWITCH on real peers runs through the NVSwitch: both multicast mappings or neither.
static int check_switch(const ShardArgs &a) {
  if ((a.stage_mc == nullptr) != (a.var_mc == nullptr))
    return fail(ACCEL_ERR_INVALID,
                "ACCEL_OVERLAP_SWITCH needs stage_multicast AND var_multicast (or neither: the "
                "peer-load form)");
  if (a.var_bf16[0] != nullptr && a.var_mc != nullptr)
    return fail(ACCEL_ERR_INVALID, "var_bf16 and the multicast pull are mutually exclusive");
  return ACCEL_OK;
}

// Share of the CTAs that reduce + apply + pull in the in-switch step, in percent. With one local
// gradient the step is a pure collective and the reducers do all the work that matters: 25 %
// (the link probe is flat from 74 CTAs up and slower at 37). With several local gradients the
// HBM-bound summers need the CTAs: 8 % (6-12 % within noise of each other at 8 GPUs, 25 % slower,
// 3 % starves the links; profiles/r02_sharded_step.md).
// ACCEL_OVERLAP_ROLES: share of the CTAs that apply + pull (the rest push), in percent.
static int roles_percent(const Tuning &t) { return t.reducer_pct > 0 ? t.reducer_pct : 50; }

static int reducer_percent(const Tuning &t, int n_workers) {
  if (t.reducer_pct > 0) return t.reducer_pct;
  return n_workers == 1 ? 25 : 8;
}

extern "C" {

int accel_shard_step_fused(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                           float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                           int n_contrib, const accel_hparams *hp, const float *grad_sumsq_dev,
                           int n_slices, int overlap, accel_stream_t stream) {
  if (!peers) return fail(ACCEL_ERR_INVALID, "peers is NULL");
  if (!overlap_ok(overlap)) return fail(ACCEL_ERR_INVALID, "unknown overlap %d", overlap);
  TaskState *ts = nullptr;
  int rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  FusedArgs<1> fa;
  rc = fill_rank(peers, optimizer, slot0, slot1, grads, n_workers, grad_dtype, n_contrib, hp,
                 grad_sumsq_dev, &fa.r[0]);
  if (rc) return rc;
  fa.n_slices = clamp_slices(n_slices, peers->shard_elems);
  fa.roles = overlap == ACCEL_OVERLAP_ROLES ? roles_percent(ctx->tuning) : 0;
  fa.n_reducers = reducer_percent(ctx->tuning, n_workers);   // percent; the launcher turns it into CTAs
  fa.trace = ctx->tuning.trace;
  fa.l2hint = ctx->tuning.l2hint;
  fa.epoch_base = ts->shard_epoch;
  fa.ctas_per_rank = 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (overlap == ACCEL_OVERLAP_SWITCH) {
    rc = check_switch(fa.r[0]);
    if (rc) return rc;
    rc = launch_switch(fa, optimizer, grad_dtype, *ts, ctx->tuning, s);
  } else if (fa.r[0].in_place) {
    return fail(ACCEL_ERR_INVALID,
                "the gradient is this rank's stage buffer: peers overwrite it during a unicast push; "
                "only ACCEL_OVERLAP_SWITCH runs in place");
  } else {
    rc = launch_fused<1>(fa, 1, optimizer, grad_dtype, *ts, ctx->tuning, s);
  }
  // the epoch advances only once the step is really enqueued: a failed launch leaves this
  // rank in step with its peers
  if (rc == ACCEL_OK) ts->shard_epoch += fa.n_slices;
  return rc;
}

int accel_shard_step_local_group(accel_ctx *ctx, const accel_peers *const *peers_of_rank,
                                 int n_ranks, int optimizer, float *const *slot0,
                                 float *const *slot1, const void *const *const *grads,
                                 int n_workers, int grad_dtype, int n_contrib,
                                 const accel_hparams *hp, const float *grad_sumsq_dev,
                                 int n_slices, int overlap, accel_stream_t stream) {
  if (!overlap_ok(overlap)) return fail(ACCEL_ERR_INVALID, "unknown overlap %d", overlap);
  if (!peers_of_rank || !grads) return fail(ACCEL_ERR_INVALID, "peers_of_rank/grads is NULL");
  if (n_ranks < 1 || n_ranks > kEmuRanks)
    return fail(ACCEL_ERR_INVALID, "n_ranks %d out of range [1,%d]", n_ranks, kEmuRanks);
  FusedArgs<kEmuRanks> fa;
  memset(&fa, 0, sizeof(fa));
  TaskState *ts0 = nullptr;
  for (int r = 0; r < n_ranks; ++r) {
    if (!peers_of_rank[r]) return fail(ACCEL_ERR_INVALID, "peers_of_rank[%d] is NULL", r);
    if (peers_of_rank[r]->rank != r || peers_of_rank[r]->world != n_ranks)
      return fail(ACCEL_ERR_INVALID, "peers_of_rank[%d] must describe rank %d of %d", r, r, n_ranks);
    TaskState *ts = nullptr;
    int rc = task_state(ctx, r, &ts);
    if (rc) return rc;
    if (r == 0) ts0 = ts;
    if (ts->device != ts0->device)
      return fail(ACCEL_ERR_INVALID, "a local group lives on one GPU; rank %d is on another", r);
    if (ts->shard_epoch != ts0->shard_epoch)
      return fail(ACCEL_ERR_STATE, "ranks of a local group must be in the same epoch");
    rc = fill_rank(peers_of_rank[r], optimizer, slot0 ? slot0[r] : nullptr,
                   slot1 ? slot1[r] : nullptr, grads[r], n_workers, grad_dtype, n_contrib, hp,
                   grad_sumsq_dev, &fa.r[r]);
    if (rc) return rc;
    if (overlap != ACCEL_OVERLAP_SWITCH && fa.r[r].in_place)
      return fail(ACCEL_ERR_INVALID, "only ACCEL_OVERLAP_SWITCH runs in place (rank %d)", r);
    if (overlap == ACCEL_OVERLAP_SWITCH && (fa.r[r].stage_mc || fa.r[r].var_mc))
      return fail(ACCEL_ERR_INVALID,
                  "a local group has no NVSwitch between its ranks: multicast pointers must be NULL");
  }
  fa.n_slices = clamp_slices(n_slices, peers_of_rank[0]->shard_elems);
  fa.roles = overlap == ACCEL_OVERLAP_ROLES ? roles_percent(ctx->tuning) : 0;
  fa.n_reducers = reducer_percent(ctx->tuning, n_workers);
  fa.trace = 0;
  fa.l2hint = 0;
  fa.epoch_base = ts0->shard_epoch;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int rc = overlap == ACCEL_OVERLAP_SWITCH
                     ? launch_switch_group(fa, n_ranks, optimizer, grad_dtype, *ts0, ctx->tuning, s)
                     : launch_fused<kEmuRanks>(fa, n_ranks, optimizer, grad_dtype, *ts0,
                                               ctx->tuning, s);
  if (rc == ACCEL_OK)
    for (int r = 0; r < n_ranks; ++r) ctx->task[r].shard_epoch += fa.n_slices;
  return rc;
}

int accel_shard_reset(accel_ctx *ctx, const accel_peers *peers, accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, 0, 0, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  ACCEL_CUDA(cudaStreamSynchronize(s));
  if (ts->shard_aux) ACCEL_CUDA(cudaStreamSynchronize(ts->shard_aux));
  ACCEL_CUDA(cudaMemsetAsync(a.flags[a.rank], 0, ACCEL_SHARD_FLAG_WORDS * sizeof(uint32_t), s));
  ACCEL_CUDA(cudaStreamSynchronize(s));
  ts->shard_epoch = 0;
  return ACCEL_OK;
}

int accel_ctx_set_step_shape(accel_ctx *ctx, int reducer_pct, int coop) {
  if (!ctx) return fail(ACCEL_ERR_INVALID, "ctx is NULL");
  if (reducer_pct < 0 || reducer_pct > 90)
    return fail(ACCEL_ERR_INVALID, "reducer_pct %d out of range [0,90]", reducer_pct);
  ctx->tuning.reducer_pct = reducer_pct;
  ctx->tuning.coop = coop ? 1 : 0;
  return ACCEL_OK;
}

}  // extern "C"
