// Fused N-way gradient aggregation + optimizer apply, vector-load variant.
//
// One pass over HBM: every thread keeps NW gradient vectors (+ the state
// vectors) of U independent items in flight, adds the gradients in worker order
// in registers, applies the update and writes the state back. A CTA walks the
// flat arena in tiles of blockDim*U items, so each of the NW+state streams is
// read in contiguous blockDim*U*VE*4-byte runs (DRAM page locality across the
// concurrent streams), and the grid is persistent (SMs x resident CTAs).
//
// Algorithmic bytes/element: NW*sizeof(G) + 8*(1 + slots)   (SURVEY §8 d).
#pragma once

#include "ps_math.cuh"

namespace accel {

// NW > 0: compile-time worker count; NW == 0: runtime a.n_workers (any count up
// to ACCEL_MAX_WORKERS), gradients fetched four workers at a time.
template <int OPT, typename G, int NW, bool WIDE, int U>
__global__ void __launch_bounds__(512)
    ps_apply_ldg(const __grid_constant__ ApplyArgs a) {
  constexpr int VE = ApplyItem<G, WIDE, OPT>::VE;
  constexpr int R = ApplyItem<G, WIDE, OPT>::R;
  constexpr int SLOTS = OptTraits<OPT>::kSlots;
  const Scalars sc = make_scalars(a);

  const uint64_t n_items = a.count / VE;
  const uint64_t tile = static_cast<uint64_t>(blockDim.x) * U;
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * tile;

  for (uint64_t base = static_cast<uint64_t>(blockIdx.x) * tile; base < n_items; base += stride) {
    const bool full = base + tile <= n_items;
    uint64_t idx[U];
    bool ok[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      idx[j] = base + static_cast<uint64_t>(j) * blockDim.x + threadIdx.x;
      ok[j] = full || idx[j] < n_items;
    }

    float acc[U][VE];
    if constexpr (NW > 0) {
      uint32_t g[NW][U][R];   // packed: bf16 gradients are widened only when they are added
#pragma unroll
      for (int w = 0; w < NW; ++w)
#pragma unroll
        for (int j = 0; j < U; ++j)
          if (ok[j]) load_grad_raw<R>(a.grad[w], idx[j], g[w][j]);
      float v[U][VE], s0[U][VE], s1[U][VE];
#pragma unroll
      for (int j = 0; j < U; ++j) {
        if (ok[j]) {
          load_state_item<VE, WIDE>(a.var, idx[j], v[j]);
          if constexpr (SLOTS >= 1) load_state_item<VE, WIDE>(a.slot0, idx[j], s0[j]);
          if constexpr (SLOTS >= 2) load_state_item<VE, WIDE>(a.slot1, idx[j], s1[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < U; ++j) {
        if (!ok[j]) continue;
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          float sum = grad_elem<G>(g[0][j], e);
#pragma unroll
          for (int w = 1; w < NW; ++w) sum = __fadd_rn(sum, grad_elem<G>(g[w][j], e));
          update<OPT>(v[j][e], s0[j][e], s1[j][e], finish_grad(sum, sc), sc);
        }
        store_state_item<VE, WIDE>(a.var, idx[j], v[j]);
        if constexpr (SLOTS >= 1) store_state_item<VE, WIDE>(a.slot0, idx[j], s0[j]);
        if constexpr (SLOTS >= 2) store_state_item<VE, WIDE>(a.slot1, idx[j], s1[j]);
      }
    } else {
      // runtime worker count: state first, then gradients 4 workers per batch
      float v[U][VE], s0[U][VE], s1[U][VE];
#pragma unroll
      for (int j = 0; j < U; ++j) {
        if (ok[j]) {
          load_state_item<VE, WIDE>(a.var, idx[j], v[j]);
          if constexpr (SLOTS >= 1) load_state_item<VE, WIDE>(a.slot0, idx[j], s0[j]);
          if constexpr (SLOTS >= 2) load_state_item<VE, WIDE>(a.slot1, idx[j], s1[j]);
        }
      }
      const int n = a.n_workers;
      for (int w0 = 0; w0 < n; w0 += 4) {
        uint32_t g[4][U][R];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (w0 + k < n)
#pragma unroll
            for (int j = 0; j < U; ++j)
              if (ok[j]) load_grad_raw<R>(a.grad[w0 + k], idx[j], g[k][j]);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (w0 + k < n)
#pragma unroll
            for (int j = 0; j < U; ++j)
#pragma unroll
              for (int e = 0; e < VE; ++e)
                acc[j][e] = (w0 + k == 0) ? grad_elem<G>(g[k][j], e)
                                          : __fadd_rn(acc[j][e], grad_elem<G>(g[k][j], e));
      }
#pragma unroll
      for (int j = 0; j < U; ++j) {
        if (!ok[j]) continue;
#pragma unroll
        for (int e = 0; e < VE; ++e)
          update<OPT>(v[j][e], s0[j][e], s1[j][e], finish_grad(acc[j][e], sc), sc);
        store_state_item<VE, WIDE>(a.var, idx[j], v[j]);
        if constexpr (SLOTS >= 1) store_state_item<VE, WIDE>(a.slot0, idx[j], s0[j]);
        if constexpr (SLOTS >= 2) store_state_item<VE, WIDE>(a.slot1, idx[j], s1[j]);
      }
    }
  }

  // scalar tail: the last count % VE elements, one thread each
  const uint64_t tail0 = n_items * VE;
  if (blockIdx.x == 0 && tail0 + threadIdx.x < a.count) {
    const uint64_t i = tail0 + threadIdx.x;
    const int n = NW > 0 ? NW : a.n_workers;
    float sum = load_grad_scalar<G>(a.grad[0], i);
    for (int w = 1; w < n; ++w) sum = __fadd_rn(sum, load_grad_scalar<G>(a.grad[w], i));
    float v = a.var[i], s0 = 0.f, s1 = 0.f;
    if constexpr (SLOTS >= 1) s0 = a.slot0[i];
    if constexpr (SLOTS >= 2) s1 = a.slot1[i];
    update<OPT>(v, s0, s1, finish_grad(sum, sc), sc);
    a.var[i] = v;
    if constexpr (SLOTS >= 1) a.slot0[i] = s0;
    if constexpr (SLOTS >= 2) a.slot1[i] = s1;
  }
}

// Launch one instantiation; returns false when (NW, WIDE, U) is not built.
template <int OPT, typename G>
int launch_apply_ldg(const ApplyArgs &a, int n_workers, bool wide, const LaunchShape &ls,
                     cudaStream_t stream);

}  // namespace accel
