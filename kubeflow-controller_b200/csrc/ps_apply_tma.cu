// Fused N-way gradient aggregation + optimizer apply, bulk-async (TMA) staged variant.
//
// One persistent CTA per SM. Warp 0 is the producer: one elected lane issues, per tile, one
// 1-D bulk copy (cp.async.bulk global -> shared, SASS UBLKCP) per worker gradient and per state
// array into a ring of kStages shared-memory stages, each guarded by a full/empty mbarrier
// pair (expect_tx / complete_tx byte counting). The eight consumer warps wait on `full`, sum
// the gradients from shared memory in worker order, apply the update in registers, store the
// state with 128-bit global stores, and release the stage on `empty`. The copy engine keeps
// kStages * (NW + state) * 4 KiB per SM in flight without spending registers or LSU issue
// slots on loads.
//
// Covers whole tiles of kTile elements; the launcher hands the remainder (< kTile elements)
// to the vector-load kernel. Arithmetic is the shared update<OPT> (bit-identical results).
#include <algorithm>

#include "ps_apply_ldg.cuh"
#include "ps_apply_tma.cuh"

namespace accel {

constexpr int kTile = 1024;            // elements per stream per stage
constexpr int kConsumerWarps = 8;      // 256 consumer threads x one float4 = kTile
constexpr int kTmaThreads = 32 * (1 + kConsumerWarps);
constexpr int kMaxStages = 6;
constexpr size_t kSmemBudget = 200 * 1024;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk copy global -> this CTA's shared memory, completion counted on `bar`.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, uint32_t bytes,
                                         uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

template <int OPT, typename G, int NW>
__global__ void __launch_bounds__(kTmaThreads, 1)
    ps_apply_tma(const __grid_constant__ ApplyArgs a, const uint64_t n_tiles, const int stages) {
  constexpr int SLOTS = OptTraits<OPT>::kSlots;
  constexpr uint32_t kGradBytes = kTile * sizeof(G);
  constexpr uint32_t kStateBytes = kTile * 4;
  constexpr uint32_t kStageBytes = NW * kGradBytes + (1 + SLOTS) * kStateBytes;

  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);                 // the producer's expect_tx arrive
      mbar_init(&empty_bar[s], kConsumerWarps);   // one arrive per consumer warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == 0) {
    // ---- producer -------------------------------------------------------------------------
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        mbar_wait(&empty_bar[s], phase ^ 1);      // first pass over the ring returns at once
        unsigned char *stage = smem + static_cast<size_t>(s) * kStageBytes;
        mbar_expect_tx(&full_bar[s], kStageBytes);
        const uint64_t e0 = t * kTile;
#pragma unroll
        for (int w = 0; w < NW; ++w)
          bulk_g2s(stage + w * kGradBytes, static_cast<const G *>(a.grad[w]) + e0, kGradBytes,
                   &full_bar[s]);
        unsigned char *st = stage + NW * kGradBytes;
        bulk_g2s(st, a.var + e0, kStateBytes, &full_bar[s]);
        if constexpr (SLOTS >= 1) bulk_g2s(st + kStateBytes, a.slot0 + e0, kStateBytes, &full_bar[s]);
        if constexpr (SLOTS >= 2)
          bulk_g2s(st + 2 * kStateBytes, a.slot1 + e0, kStateBytes, &full_bar[s]);
        if (++s == stages) {
          s = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ---- consumers ------------------------------------------------------------------------
    const Scalars sc = make_scalars(a);
    const int ct = threadIdx.x - 32;              // 0..255: owns elements 4*ct .. 4*ct+3 of a tile
    int s = 0;
    uint32_t phase = 0;
    for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      mbar_wait(&full_bar[s], phase);
      const unsigned char *stage = smem + static_cast<size_t>(s) * kStageBytes;
      float g[NW][4];
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        if constexpr (sizeof(G) == 4) {
          const float4 x = *reinterpret_cast<const float4 *>(stage + w * kGradBytes + ct * 16);
          g[w][0] = x.x, g[w][1] = x.y, g[w][2] = x.z, g[w][3] = x.w;
        } else {
          const uint2 x = *reinterpret_cast<const uint2 *>(stage + w * kGradBytes + ct * 8);
          g[w][0] = bf16_lo(x.x), g[w][1] = bf16_hi(x.x);
          g[w][2] = bf16_lo(x.y), g[w][3] = bf16_hi(x.y);
        }
      }
      const unsigned char *st = stage + NW * kGradBytes;
      float v[4], s0[4], s1[4];
      *reinterpret_cast<float4 *>(v) = *reinterpret_cast<const float4 *>(st + ct * 16);
      if constexpr (SLOTS >= 1)
        *reinterpret_cast<float4 *>(s0) =
            *reinterpret_cast<const float4 *>(st + kStateBytes + ct * 16);
      if constexpr (SLOTS >= 2)
        *reinterpret_cast<float4 *>(s1) =
            *reinterpret_cast<const float4 *>(st + 2 * kStateBytes + ct * 16);
      // everything this thread needs from the stage is in registers: hand the stage back
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float sum = g[0][e];
#pragma unroll
        for (int w = 1; w < NW; ++w) sum = __fadd_rn(sum, g[w][e]);
        update<OPT>(v[e], s0[e], s1[e], finish_grad(sum, sc), sc);
      }
      const uint64_t e0 = t * kTile + static_cast<uint64_t>(ct) * 4;
      st_state_128(a.var + e0, v);
      if constexpr (SLOTS >= 1) st_state_128(a.slot0 + e0, s0);
      if constexpr (SLOTS >= 2) st_state_128(a.slot1 + e0, s1);
      if (++s == stages) {
        s = 0;
        phase ^= 1;
      }
    }
  }
}

template <int OPT, typename G, int NW>
static int launch_tma_one(const ApplyArgs &a, uint64_t n_tiles, const TaskState &ts,
                          const Tuning &tune, cudaStream_t stream) {
  constexpr size_t kStageBytes =
      NW * kTile * sizeof(G) + (1 + OptTraits<OPT>::kSlots) * kTile * 4;
  int stages = static_cast<int>(kSmemBudget / kStageBytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (tune.unroll) stages = std::min(stages, std::max(2, tune.unroll * 2));  // sweep knob: 2/4/6(8)
  if (stages < 2) return ACCEL_ERR_STATE;
  auto kern = ps_apply_tma<OPT, G, NW>;
  const size_t smem = kStageBytes * stages;
  ACCEL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(smem)));
  int grid = ts.sm_count;
  if (static_cast<uint64_t>(grid) > n_tiles) grid = static_cast<int>(n_tiles);
  kern<<<grid, kTmaThreads, smem, stream>>>(a, n_tiles, stages);
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}

template <int OPT, typename G>
static int launch_tma_nw(const ApplyArgs &a, uint64_t n_tiles, const TaskState &ts,
                         const Tuning &tune, cudaStream_t stream) {
  switch (a.n_workers) {
    case 1: return launch_tma_one<OPT, G, 1>(a, n_tiles, ts, tune, stream);
    case 2: return launch_tma_one<OPT, G, 2>(a, n_tiles, ts, tune, stream);
    case 4: return launch_tma_one<OPT, G, 4>(a, n_tiles, ts, tune, stream);
    case 8: return launch_tma_one<OPT, G, 8>(a, n_tiles, ts, tune, stream);
    default: return ACCEL_ERR_STATE;   // other worker counts: vector-load variant
  }
}

int launch_apply_tma(int optimizer, int grad_dtype, const ApplyArgs &a, const TaskState &ts,
                     const Tuning &tune, cudaStream_t stream) {
  const uint64_t n_tiles = a.count / kTile;
  if (n_tiles == 0) return ACCEL_ERR_STATE;
  const bool bf = grad_dtype == ACCEL_BF16;
  int rc;
  switch (optimizer) {
    case ACCEL_OPT_SGD:
      rc = bf ? launch_tma_nw<OPT_SGD, bf16_t>(a, n_tiles, ts, tune, stream)
              : launch_tma_nw<OPT_SGD, float>(a, n_tiles, ts, tune, stream);
      break;
    case ACCEL_OPT_MOMENTUM:
      rc = bf ? launch_tma_nw<OPT_MOMENTUM, bf16_t>(a, n_tiles, ts, tune, stream)
              : launch_tma_nw<OPT_MOMENTUM, float>(a, n_tiles, ts, tune, stream);
      break;
    case ACCEL_OPT_ADAM:
      rc = bf ? launch_tma_nw<OPT_ADAM, bf16_t>(a, n_tiles, ts, tune, stream)
              : launch_tma_nw<OPT_ADAM, float>(a, n_tiles, ts, tune, stream);
      break;
    default:
      return fail(ACCEL_ERR_INVALID, "unknown optimizer %d", optimizer);
  }
  if (rc != ACCEL_OK) return rc;
  // remainder (< kTile elements): the vector-load kernel on the tail of every array
  const uint64_t done = n_tiles * kTile;
  if (done == a.count) return ACCEL_OK;
  ApplyArgs tail = a;
  tail.count = a.count - done;
  tail.var += done;
  if (tail.slot0) tail.slot0 += done;
  if (tail.slot1) tail.slot1 += done;
  const size_t gsize = bf ? 2 : 4;
  for (int w = 0; w < a.n_workers; ++w)
    tail.grad[w] = static_cast<const char *>(a.grad[w]) + done * gsize;
  Tuning ldg;
  ldg.kernel = ACCEL_KERNEL_LDG128;
  return launch_apply(optimizer, grad_dtype, tail, ts, ldg, stream);
}

}  // namespace accel
