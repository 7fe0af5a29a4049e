// Dispatcher from runtime (n_workers, width, unroll) to the compiled
// instantiations of ps_apply_ldg. Included by one small .cu per (optimizer,
// gradient dtype) so the instantiations compile in parallel.
#pragma once

#include <algorithm>

#include "ps_apply_ldg.cuh"

namespace accel {

// Registers are the budget: keep (packed gradient + state) registers in flight per thread
// at or below ~92 so nothing spills under __launch_bounds__(512) (128 regs).
constexpr int max_unroll(int nw_in_flight, int r, int ve, int slots) {
  int per_u = nw_in_flight * r + (1 + slots) * ve;
  int u = 92 / per_u;
  return u >= 4 ? 4 : (u >= 2 ? 2 : 1);
}

template <int OPT, typename G, int NW, bool WIDE, int U>
static int launch_one(const ApplyArgs &a, const LaunchShape &ls, cudaStream_t stream) {
  constexpr int VE = ApplyItem<G, WIDE, OPT>::VE;
  constexpr int kMaxU =
      max_unroll(NW > 0 ? NW : 6, ApplyItem<G, WIDE, OPT>::R, VE, OptTraits<OPT>::kSlots);
  if constexpr (U > kMaxU) {
    return launch_one<OPT, G, NW, WIDE, (U > 1 ? U / 2 : 1)>(a, ls, stream);
  } else {
    auto kern = ps_apply_ldg<OPT, G, NW, WIDE, U>;
    int grid = ls.grid;
    if (grid <= 0) {
      int per_sm = 0, dev = 0, sms = 0;
      ACCEL_CUDA(cudaGetDevice(&dev));
      ACCEL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      ACCEL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, ls.threads, 0));
      if (per_sm < 1) per_sm = 1;
      // With eight or more concurrent streams two resident CTAs already keep > 100 KB per SM in
      // flight; a third only adds DRAM page conflicts (bf16 Adam, N=8: 0.644 ms at 3, 0.628 at 2).
      constexpr int kStreams = (NW > 0 ? NW : 6) + 1 + OptTraits<OPT>::kSlots;
      if (ls.grid == 0 && kStreams >= 8) per_sm = std::min(per_sm, 2);
      if (ls.grid < 0) per_sm = std::min(per_sm, -ls.grid);  // -k = at most k CTAs per SM
      grid = per_sm * sms;
    }
    const uint64_t items = a.count / VE;
    const uint64_t tile = static_cast<uint64_t>(ls.threads) * U;
    const uint64_t tiles = std::max<uint64_t>(1, (items + tile - 1) / tile);
    if (static_cast<uint64_t>(grid) > tiles) grid = static_cast<int>(tiles);
    kern<<<grid, ls.threads, 0, stream>>>(a);
    ACCEL_CUDA(cudaGetLastError());
    return ACCEL_OK;
  }
}

template <int OPT, typename G, int NW, bool WIDE>
static int launch_u(const ApplyArgs &a, const LaunchShape &ls, cudaStream_t stream) {
  switch (ls.unroll) {
    case 1:
      return launch_one<OPT, G, NW, WIDE, 1>(a, ls, stream);
    case 2:
      return launch_one<OPT, G, NW, WIDE, 2>(a, ls, stream);
    default:
      return launch_one<OPT, G, NW, WIDE, 4>(a, ls, stream);
  }
}

template <int OPT, typename G, bool WIDE>
static int launch_nw(const ApplyArgs &a, int n_workers, const LaunchShape &ls,
                     cudaStream_t stream) {
  switch (n_workers) {
    case 1:
      return launch_u<OPT, G, 1, WIDE>(a, ls, stream);
    case 2:
      return launch_u<OPT, G, 2, WIDE>(a, ls, stream);
    case 4:
      return launch_u<OPT, G, 4, WIDE>(a, ls, stream);
    case 8:
      return launch_u<OPT, G, 8, WIDE>(a, ls, stream);
    default:
      return launch_u<OPT, G, 0, WIDE>(a, ls, stream);
  }
}

template <int OPT, typename G>
int launch_apply_ldg(const ApplyArgs &a, int n_workers, bool wide, const LaunchShape &ls,
                     cudaStream_t stream) {
  return wide ? launch_nw<OPT, G, true>(a, n_workers, ls, stream)
              : launch_nw<OPT, G, false>(a, n_workers, ls, stream);
}

}  // namespace accel

#define ACCEL_INSTANTIATE_APPLY_LDG(OPT, G)                                              \
  template int accel::launch_apply_ldg<OPT, G>(const accel::ApplyArgs &, int, bool,      \
                                               const accel::LaunchShape &, cudaStream_t);
