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
#include "ps_shard.cuh"

namespace accel {

// ---- push: local N-way sum fused with the reduce-scatter -----------------------
// Items of 8 elements, handed out in granules of kPushGranule items. Granule g of a rank goes
// to destination (rank + 1 + g) mod world: consecutive granules go to consecutive owners, so
// through the whole kernel every rank is storing to ALL owners at once (its own included) —
// NVLink egress and ingress are busy from the first tile to the last, and no owner is a hot
// spot at any moment.
template <typename G, int NW, int U>
__global__ void __launch_bounds__(512) ps_push(const __grid_constant__ ShardArgs a) {
  const uint64_t shard_items = a.shard / 8;
  const uint64_t sub_items = a.sub_count / 8;   // items per shard in this launch
  const uint64_t sub_first = a.sub_begin / 8;
  const uint64_t granules = (sub_items + kPushGranule - 1) / kPushGranule;   // per owner
  const uint64_t n_items = granules * kPushGranule * a.world;
  const uint64_t tile = static_cast<uint64_t>(blockDim.x) * U;
  const int n = NW > 0 ? NW : a.n_workers;

  for (uint64_t base = static_cast<uint64_t>(blockIdx.x) * tile; base < n_items;
       base += static_cast<uint64_t>(gridDim.x) * tile) {
    uint64_t item[U];   // arena item (8 elements) this thread sums
    uint64_t dst[U];    // item index inside the owner's stage slot [rank]
    int owner[U];
    bool ok[U];
    float acc[U][8];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const uint64_t l = base + static_cast<uint64_t>(j) * blockDim.x + threadIdx.x;
      const uint64_t g = l / kPushGranule;
      const uint64_t in_shard = (g / a.world) * kPushGranule + (l % kPushGranule);
      ok[j] = l < n_items && in_shard < sub_items;
      int c = a.rank + 1 + static_cast<int>(g % a.world);
      if (c >= a.world) c -= a.world;
      owner[j] = ok[j] ? c : a.rank;
      const uint64_t within = sub_first + (ok[j] ? in_shard : 0);
      item[j] = static_cast<uint64_t>(owner[j]) * shard_items + within;
      dst[j] = static_cast<uint64_t>(a.rank) * shard_items + within;
    }
    if constexpr (sizeof(G) == 2 && NW == 1) {
      if (a.stage_bf16) {
        // one bf16 gradient per rank: the partial IS the gradient — forward the raw item
        uint32_t raw[U][4];
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j]) ldg_stream_128u(static_cast<const uint16_t *>(a.grad[0]) + item[j] * 8, raw[j]);
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j])
            st_128u(reinterpret_cast<uint16_t *>(a.stage[owner[j]]) + dst[j] * 8, raw[j]);
        continue;
      }
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
    for (int j = 0; j < U; ++j)
      if (ok[j]) st_256(a.stage[owner[j]] + dst[j] * 8, acc[j]);
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
  const Scalars sc = make_scalars(sa);

  const uint64_t shard_items = a.shard / 8;
  const uint64_t sub_first = a.sub_begin / 8, sub_end = sub_first + a.sub_count / 8;
  const float *stage = a.stage[a.rank];
  const uint64_t var_off = static_cast<uint64_t>(a.rank) * a.shard;
  const uint64_t tile = static_cast<uint64_t>(blockDim.x) * U;

  for (uint64_t base = sub_first + static_cast<uint64_t>(blockIdx.x) * tile; base < sub_end;
       base += static_cast<uint64_t>(gridDim.x) * tile) {
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const uint64_t it = base + static_cast<uint64_t>(j) * blockDim.x + threadIdx.x;
      if (it >= sub_end) continue;
      float acc[8], v[8], s0[8], s1[8];
      ld_state_256(a.var[a.rank] + var_off + it * 8, v);
      if constexpr (SLOTS >= 1) ld_state_256(a.slot0 + it * 8, s0);
      if constexpr (SLOTS >= 2) ld_state_256(a.slot1 + it * 8, s1);
      // stage slots in rank order, four loads in flight at a time
      for (int s0i = 0; s0i < a.world; s0i += 4) {
        float g[4][8];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (s0i + k < a.world) ld_stage_item(a, stage, s0i + k, shard_items, it, g[k]);
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
      pull_item(a, var_off + it * 8, v);
    }
  }
  signal_all(a, kFlagTicketApply, kFlagPull);
}

__global__ void ps_wait_pull(const __grid_constant__ ShardArgs a) {
  wait_all(a.flags[a.rank], kFlagPull, a.world, a.epoch);
}

__global__ void ps_wait_push(const __grid_constant__ ShardArgs a) {
  wait_all(a.flags[a.rank], kFlagPush, a.world, a.epoch);
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

int fill_peers(const accel_peers *peers, uint64_t sub_begin, uint64_t sub_count, ShardArgs *a) {
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
  if (sub_count == 0) {  // the whole shard
    sub_begin = 0;
    sub_count = peers->shard_elems;
  }
  if (sub_begin % 8 || sub_count % 8 || sub_begin + sub_count > peers->shard_elems)
    return fail(ACCEL_ERR_INVALID,
                "sub-range [%llu, +%llu) must be multiples of 8 inside the shard of %llu elements",
                static_cast<unsigned long long>(sub_begin),
                static_cast<unsigned long long>(sub_count),
                static_cast<unsigned long long>(peers->shard_elems));
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
  if ((reinterpret_cast<uintptr_t>(peers->var_multicast) |
       reinterpret_cast<uintptr_t>(peers->stage_multicast)) & 31)
    return fail(ACCEL_ERR_ALIGN, "var_multicast / stage_multicast must be 32-byte aligned");
  a->var_mc = peers->var_multicast;
  a->stage_mc = peers->stage_multicast;
  if (peers->stage_dtype != ACCEL_F32 && peers->stage_dtype != ACCEL_BF16)
    return fail(ACCEL_ERR_INVALID, "unknown stage_dtype %d", peers->stage_dtype);
  a->stage_bf16 = peers->stage_dtype == ACCEL_BF16;
  if (peers->var_bf16) {
    for (int r = 0; r < peers->world; ++r) {
      if (!peers->var_bf16[r]) return fail(ACCEL_ERR_INVALID, "peer %d has a NULL var_bf16", r);
      if (reinterpret_cast<uintptr_t>(peers->var_bf16[r]) & 15)
        return fail(ACCEL_ERR_ALIGN, "peer %d var_bf16 must be 16-byte aligned", r);
      a->var_bf16[r] = peers->var_bf16[r];
    }
    if (peers->var_multicast)
      return fail(ACCEL_ERR_INVALID, "var_bf16 and var_multicast are mutually exclusive");
  }
  a->rank = peers->rank;
  a->world = peers->world;
  a->shard = peers->shard_elems;
  a->sub_begin = sub_begin;
  a->sub_count = sub_count;
  return ACCEL_OK;
}

int fill_grads(const void *const *grads, int n_workers, int grad_dtype, ShardArgs *a) {
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
    a->grad[w] = grads[w];
  }
  a->n_workers = n_workers;
  // zero-copy form of the in-switch step: the worker wrote its gradient straight into this rank's
  // stage buffer (arena layout), so there is nothing to sum or copy before the switch reads it
  a->in_place = n_workers == 1 && grads[0] == static_cast<const void *>(a->stage[a->rank]) &&
                (grad_dtype == ACCEL_F32 || a->stage_bf16);
  // any other overlap of a gradient with the stage buffer is a race (peers write into it)
  const char *lo = reinterpret_cast<const char *>(a->stage[a->rank]);
  const char *hi = lo + static_cast<size_t>(a->shard) * a->world * sizeof(float);
  for (int w = 0; w < n_workers; ++w) {
    const char *g = static_cast<const char *>(grads[w]);
    if (g >= lo && g < hi && !a->in_place)
      return fail(ACCEL_ERR_INVALID,
                  "grads[%d] lies inside this rank's stage buffer; only the in-switch step accepts "
                  "that, with n_workers = 1 and grads[0] == stage[rank]", w);
  }
  if (a->stage_bf16 && !(grad_dtype == ACCEL_BF16 && n_workers == 1))
    return fail(ACCEL_ERR_INVALID,
                "stage_dtype ACCEL_BF16 carries ONE raw bf16 gradient per rank; got n_workers %d, "
                "grad_dtype %d", n_workers, grad_dtype);
  return ACCEL_OK;
}

int fill_optimizer(int optimizer, float *slot0, float *slot1, int n_contrib,
                   const accel_hparams *hp, const float *grad_sumsq_dev, ShardArgs *a) {
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
  if (hp->aggregate != ACCEL_AGG_SUM && hp->aggregate != ACCEL_AGG_MEAN)
    return fail(ACCEL_ERR_INVALID, "unknown aggregate %d", hp->aggregate);
  if (n_contrib < 1) return fail(ACCEL_ERR_INVALID, "n_contrib must be >= 1");
  a->slot0 = slot0;
  a->slot1 = slot1;
  a->sumsq = grad_sumsq_dev;
  a->use_nesterov = hp->use_nesterov ? 1 : 0;
  a->lr = hp->lr;
  a->momentum = hp->momentum;
  a->beta1 = hp->beta1;
  a->one_minus_beta1 = 1.0f - hp->beta1;
  a->one_minus_beta2 = 1.0f - hp->beta2;
  a->epsilon = hp->epsilon;
  a->alpha = 0.0f;
  if (optimizer == ACCEL_OPT_ADAM) {
    const float num = hp->lr * sqrtf(1.0f - hp->beta2_power);
    a->alpha = num / (1.0f - hp->beta1_power);
  }
  a->mean_div = hp->aggregate == ACCEL_AGG_MEAN ? static_cast<float>(n_contrib) : 0.0f;
  a->clip_norm = hp->clip_norm > 0.0f ? hp->clip_norm : 0.0f;
  return ACCEL_OK;
}

static int grid_for(const TaskState &ts, uint64_t items, int ctas_per_sm) {
  int grid = ts.sm_count * ctas_per_sm;
  const uint64_t want = (items + 255) / 256;
  if (static_cast<uint64_t>(grid) > want) grid = static_cast<int>(want);
  return grid < 1 ? 1 : grid;
}

// ctas_per_sm: grid size in CTAs per SM. 4 measured best for the NVLink-bound cases (peer-store
// rate 650 GB/s at 2, 669 at 4, flat beyond); with 8 gradients only 2 are resident (96 registers)
// and the rest queue. The two-stream pipelined step passes 1 so that the apply_pull CTAs of the
// previous slice can be resident beside the push.
static int do_push(accel_ctx *ctx, const accel_peers *peers, const void *const *grads,
                   int n_workers, int grad_dtype, uint64_t sub_begin, uint64_t sub_count,
                   cudaStream_t s, uint32_t *epoch_out, int ctas_per_sm = 4) {
  ShardArgs a;
  int rc = fill_peers(peers, sub_begin, sub_count, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  rc = fill_grads(grads, n_workers, grad_dtype, &a);
  if (rc) return rc;
  if (a.in_place)
    return fail(ACCEL_ERR_INVALID,
                "the gradient is this rank's stage buffer: peers overwrite it during a unicast push; "
                "only ACCEL_OVERLAP_SWITCH runs in place");
  a.epoch = ts->shard_epoch + 1;  // every push opens a new epoch ...
  if (ctx->tuning.ctas_per_sm) ctas_per_sm = ctx->tuning.ctas_per_sm;   // sweep knob
  const int grid = grid_for(*ts, a.sub_count / 8 * a.world, ctas_per_sm);
  rc = grad_dtype == ACCEL_BF16 ? launch_push<bf16_t>(a, grid, s) : launch_push<float>(a, grid, s);
  if (rc) return rc;
  ts->shard_epoch = a.epoch;      // ... once its launch has succeeded: a failed launch leaves this
  if (epoch_out) *epoch_out = a.epoch;   // rank in step with its peers
  return ACCEL_OK;
}

static int do_apply_pull(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                         float *slot1, int n_contrib, const accel_hparams *hp,
                         const float *grad_sumsq_dev, uint64_t sub_begin, uint64_t sub_count,
                         uint32_t epoch, cudaStream_t s, int ctas_per_sm = 4) {
  ShardArgs a;
  int rc = fill_peers(peers, sub_begin, sub_count, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  rc = fill_optimizer(optimizer, slot0, slot1, n_contrib, hp, grad_sumsq_dev, &a);
  if (rc) return rc;
  if (ts->shard_epoch == 0)
    return fail(ACCEL_ERR_STATE, "accel_shard_apply_pull before any accel_shard_push");
  a.epoch = epoch ? epoch : ts->shard_epoch;
  if (ctx->tuning.ctas_per_sm) ctas_per_sm = ctx->tuning.ctas_per_sm;   // sweep knob
  const int grid = grid_for(*ts, a.sub_count / 8, ctas_per_sm);
  switch (optimizer) {
    case ACCEL_OPT_SGD: ps_apply_pull<OPT_SGD, 1><<<grid, 256, 0, s>>>(a); break;
    case ACCEL_OPT_MOMENTUM: ps_apply_pull<OPT_MOMENTUM, 1><<<grid, 256, 0, s>>>(a); break;
    default: ps_apply_pull<OPT_ADAM, 1><<<grid, 256, 0, s>>>(a); break;
  }
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

static int do_wait_pull(accel_ctx *ctx, const accel_peers *peers, uint32_t epoch, cudaStream_t s) {
  ShardArgs a;
  int rc = fill_peers(peers, 0, 0, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  if (ts->shard_epoch == 0)
    return fail(ACCEL_ERR_STATE, "accel_shard_wait_pull before any accel_shard_push");
  a.epoch = epoch ? epoch : ts->shard_epoch;
  ps_wait_pull<<<1, 32, 0, s>>>(a);
  ACCEL_CUDA(cudaGetLastError());
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
                     int n_workers, int grad_dtype, uint64_t sub_begin, uint64_t sub_count,
                     accel_stream_t stream, uint32_t *epoch_out) {
  return do_push(ctx, peers, grads, n_workers, grad_dtype, sub_begin, sub_count,
                 static_cast<cudaStream_t>(stream), epoch_out);
}

int accel_shard_apply_pull(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                           float *slot1, int n_contrib, const accel_hparams *hp,
                           const float *grad_sumsq_dev, uint64_t sub_begin, uint64_t sub_count,
                           uint32_t epoch, accel_stream_t stream) {
  return do_apply_pull(ctx, peers, optimizer, slot0, slot1, n_contrib, hp, grad_sumsq_dev,
                       sub_begin, sub_count, epoch, static_cast<cudaStream_t>(stream));
}

int accel_shard_wait_pull(accel_ctx *ctx, const accel_peers *peers, uint32_t epoch,
                          accel_stream_t stream) {
  return do_wait_pull(ctx, peers, epoch, static_cast<cudaStream_t>(stream));
}

int accel_shard_wait_push(accel_ctx *ctx, const accel_peers *peers, uint32_t epoch,
                          accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, 0, 0, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  if (ts->shard_epoch == 0)
    return fail(ACCEL_ERR_STATE, "accel_shard_wait_push before any accel_shard_push");
  a.epoch = epoch ? epoch : ts->shard_epoch;
  ps_wait_push<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a);
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

int accel_shard_step(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                     float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                     int n_contrib, const accel_hparams *hp, const float *grad_sumsq_dev,
                     int n_chunks, accel_stream_t stream) {
  if (!peers) return fail(ACCEL_ERR_INVALID, "peers is NULL");
  if (n_chunks < 1 || n_chunks > 64)
    return fail(ACCEL_ERR_INVALID, "n_chunks %d out of range [1,64]", n_chunks);
  TaskState *ts = nullptr;
  int rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (!ts->shard_aux) {
    ACCEL_CUDA(cudaStreamCreateWithFlags(&ts->shard_aux, cudaStreamNonBlocking));
    ACCEL_CUDA(cudaEventCreateWithFlags(&ts->shard_ev_push, cudaEventDisableTiming));
    ACCEL_CUDA(cudaEventCreateWithFlags(&ts->shard_ev_done, cudaEventDisableTiming));
  }
  // chunk = the same slice of every shard; slices are multiples of 8 elements
  const uint64_t shard = peers->shard_elems;
  uint64_t per = (shard / 8 + n_chunks - 1) / n_chunks * 8;
  if (per == 0) per = 8;
  uint32_t epoch = 0;
  for (uint64_t begin = 0; begin < shard; begin += per) {
    const uint64_t n = begin + per <= shard ? per : shard - begin;
    // main stream: HBM-bound local sum + NVLink push of slice k ...
    rc = do_push(ctx, peers, grads, n_workers, grad_dtype, begin, n, s, &epoch,
                 n_chunks > 1 ? 1 : 4);
    if (rc) return rc;
    ACCEL_CUDA(cudaEventRecord(ts->shard_ev_push, s));
    // ... while the auxiliary stream applies and broadcasts slice k-1 (NVLink-bound)
    ACCEL_CUDA(cudaStreamWaitEvent(ts->shard_aux, ts->shard_ev_push, 0));
    rc = do_apply_pull(ctx, peers, optimizer, slot0, slot1, n_contrib, hp, grad_sumsq_dev, begin,
                       n, epoch, ts->shard_aux, n_chunks > 1 ? 2 : 4);
    if (rc) return rc;
  }
  ACCEL_CUDA(cudaEventRecord(ts->shard_ev_done, ts->shard_aux));
  ACCEL_CUDA(cudaStreamWaitEvent(s, ts->shard_ev_done, 0));
  return do_wait_pull(ctx, peers, epoch, s);
}

// `rows` runs of `width` bytes, `pitch` bytes apart on both sides: one 1-D copy per run (plain
// DMA transfers of a few MB each; no pitch limit to respect for multi-GiB shards).
static int copy_rows(void *dst, const void *src, size_t pitch, size_t width, int rows,
                     cudaMemcpyKind kind, cudaStream_t s) {
  if (width == pitch) {
    ACCEL_CUDA(cudaMemcpyAsync(dst, src, pitch * rows, kind, s));
    return ACCEL_OK;
  }
  for (int r = 0; r < rows; ++r)
    ACCEL_CUDA(cudaMemcpyAsync(static_cast<char *>(dst) + pitch * r,
                               static_cast<const char *>(src) + pitch * r, width, kind, s));
  return ACCEL_OK;
}

int accel_shard_step_host(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                          float *slot1, const void *const *host_grads, int n_workers,
                          int grad_dtype, int n_contrib, const accel_hparams *hp,
                          float *host_var_out, int n_slices) {
  ShardArgs a;
  int rc = fill_peers(peers, 0, 0, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  if (!hp) return fail(ACCEL_ERR_INVALID, "hparams is NULL");
  if (hp->clip_norm > 0.0f)
    return fail(ACCEL_ERR_INVALID,
                "accel_shard_step_host streams the gradient in slices and cannot clip by global "
                "norm; clip_norm must be 0");
  if (!host_grads) return fail(ACCEL_ERR_INVALID, "host_grads is NULL");
  if (n_workers < 1 || n_workers > ACCEL_MAX_WORKERS)
    return fail(ACCEL_ERR_INVALID, "n_workers %d out of range [1,%d]", n_workers,
                ACCEL_MAX_WORKERS);
  if (grad_dtype != ACCEL_F32 && grad_dtype != ACCEL_BF16)
    return fail(ACCEL_ERR_INVALID, "unknown grad_dtype %d", grad_dtype);
  for (int w = 0; w < n_workers; ++w)
    if (!host_grads[w]) return fail(ACCEL_ERR_INVALID, "host_grads[%d] is NULL", w);
  if (n_slices < 1 || n_slices > TaskState::kHostSlices)
    return fail(ACCEL_ERR_INVALID, "n_slices %d out of range [1,%d]", n_slices,
                TaskState::kHostSlices);
  const size_t gsize = grad_dtype == ACCEL_BF16 ? 2 : 4;
  const uint64_t shard = peers->shard_elems;
  const int world = peers->world;
  const size_t arena_bytes = (static_cast<size_t>(shard) * world * gsize + 255) & ~size_t{255};
  rc = ensure_host_streams(*ts);
  if (rc) return rc;
  if (ts->host_step_grad_bytes < arena_bytes * n_workers) {
    if (ts->host_step_grads) ACCEL_CUDA(cudaFree(ts->host_step_grads));
    ts->host_step_grads = nullptr;
    ts->host_step_grad_bytes = 0;
    ACCEL_CUDA(cudaMalloc(&ts->host_step_grads, arena_bytes * n_workers));
    ts->host_step_grad_bytes = arena_bytes * n_workers;
  }
  for (int k = 0; k < n_slices; ++k) {
    if (!ts->ev_slice_in[k])
      ACCEL_CUDA(cudaEventCreateWithFlags(&ts->ev_slice_in[k], cudaEventDisableTiming));
    if (!ts->ev_slice_done[k])
      ACCEL_CUDA(cudaEventCreateWithFlags(&ts->ev_slice_done[k], cudaEventDisableTiming));
  }
  const void *dev_grads[ACCEL_MAX_WORKERS];
  for (int w = 0; w < n_workers; ++w)
    dev_grads[w] = static_cast<char *>(ts->host_step_grads) + arena_bytes * w;
  uint64_t per = (shard / 8 + n_slices - 1) / n_slices * 8;
  if (per == 0) per = 8;

  auto enqueue = [&]() -> int {
    uint32_t epoch = 0;
    int k = 0;
    for (uint64_t begin = 0; begin < shard; begin += per, ++k) {
      const uint64_t n = begin + per <= shard ? per : shard - begin;
      // in: this slice of every shard of every local worker's gradient, over PCIe
      for (int w = 0; w < n_workers; ++w) {
        int rc2 = copy_rows(static_cast<char *>(const_cast<void *>(dev_grads[w])) + begin * gsize,
                            static_cast<const char *>(host_grads[w]) + begin * gsize,
                            shard * gsize, n * gsize, world, cudaMemcpyHostToDevice, ts->copy_in);
        if (rc2) return rc2;
      }
      ACCEL_CUDA(cudaEventRecord(ts->ev_slice_in[k], ts->copy_in));
      ACCEL_CUDA(cudaStreamWaitEvent(ts->compute, ts->ev_slice_in[k], 0));
      // the step on this slice: push (sum + reduce-scatter), apply + pull (all-gather)
      int rc2 = do_push(ctx, peers, dev_grads, n_workers, grad_dtype, begin, n, ts->compute, &epoch);
      if (rc2) return rc2;
      rc2 = do_apply_pull(ctx, peers, optimizer, slot0, slot1, n_contrib, hp, nullptr, begin, n,
                          epoch, ts->compute);
      if (rc2) return rc2;
      if (host_var_out) {
        // out: once every rank's pull of this slice has landed in the local replica
        rc2 = do_wait_pull(ctx, peers, epoch, ts->compute);
        if (rc2) return rc2;
        ACCEL_CUDA(cudaEventRecord(ts->ev_slice_done[k], ts->compute));
        ACCEL_CUDA(cudaStreamWaitEvent(ts->copy_out, ts->ev_slice_done[k], 0));
        rc2 = copy_rows(host_var_out + begin, a.var[a.rank] + begin, shard * sizeof(float),
                        n * sizeof(float), world, cudaMemcpyDeviceToHost, ts->copy_out);
        if (rc2) return rc2;
      }
    }
    if (!host_var_out) return do_wait_pull(ctx, peers, epoch, ts->compute);
    return ACCEL_OK;
  };
  rc = enqueue();
  const cudaError_t e1 = cudaStreamSynchronize(ts->copy_in);
  const cudaError_t e2 = cudaStreamSynchronize(ts->compute);
  const cudaError_t e3 = cudaStreamSynchronize(ts->copy_out);
  if (rc) return rc;
  if (e1 != cudaSuccess) return fail_cuda(e1, "cudaStreamSynchronize(copy_in)");
  if (e2 != cudaSuccess) return fail_cuda(e2, "cudaStreamSynchronize(compute)");
  if (e3 != cudaSuccess) return fail_cuda(e3, "cudaStreamSynchronize(copy_out)");
  uint32_t status = 0;
  ACCEL_CUDA(cudaMemcpy(&status, a.flags[a.rank] + kFlagStatus, sizeof(status),
                        cudaMemcpyDeviceToHost));
  if (status != 0)
    return fail(ACCEL_ERR_STATE, "rank %d timed out waiting for rank %u during the host-buffer step",
                a.rank, (status >> 8) & 0xff);
  return ACCEL_OK;
}

int accel_shard_check(accel_ctx *ctx, const accel_peers *peers, accel_stream_t stream) {
  ShardArgs a;
  int rc = fill_peers(peers, 0, 0, &a);
  if (rc) return rc;
  TaskState *ts = nullptr;
  rc = task_state(ctx, peers->rank, &ts);
  if (rc) return rc;
  ACCEL_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  if (ts->shard_aux) ACCEL_CUDA(cudaStreamSynchronize(ts->shard_aux));
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
