// Development probes, NOT part of libaccel_ps.so or its ABI: raw rates of the memory operations a
// fused multi-GPU kernel can use across NVLink. Built into tools/probe/libaccel_probe.so by
// kubeflow-controller_b200/build.py; tools/nvls_probe.py and tools/peer_store_probe.py drive it
// under torchrun.
//
// mode 0: multimem.st of this rank's shard through the multicast mapping (all-gather pattern)
// mode 1: multimem.ld_reduce of this rank's shard (in-switch sum over all ranks) -> `local`
// mode 2: plain local copy of the shard (HBM reference)
// mode 3: unicast peer LOAD, 128-bit   (`mc` = the peer's buffer)
// mode 4: unicast peer STORE, 128-bit
// mode 5: unicast peer STORE, 256-bit (st.global.v8.f32, what the step kernels use)
// mode 6: unicast peer STORE as bulk-async copies: each warp stages 4 KiB in shared memory
//         (256-bit loads) and one lane issues cp.async.bulk.global.shared::cta — one 4 KiB
//         request instead of 128 stores of 32 B
// mode 7: mode 1 + mode 0 in one loop (ld_reduce own shard of `mc`, multimem.st it to `mc2`): the
//         steady state of the in-switch step, both link directions busy at once
// mode 8: the same two operations DECOUPLED: even CTAs only ld_reduce the own shard of `mc` (into
//         `local`), odd CTAs only multimem.st the own shard (from `local`) to `mc2` — is the
//         fabric full-duplex when no thread waits for a reduce before it may store?
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

__device__ __forceinline__ void ld128(const float *p, float *o) {
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void st128(float *p, const float *o) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(o[0]),
               "f"(o[1]), "f"(o[2]), "f"(o[3]) : "memory");
}
__device__ __forceinline__ void ld256(const float *p, float *o) {
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]), "=f"(o[4]), "=f"(o[5]),
                 "=f"(o[6]), "=f"(o[7]) : "l"(p) : "memory");
}
__device__ __forceinline__ void st256(float *p, const float *o) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(o[0]), "f"(o[1]),
               "f"(o[2]), "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7]) : "memory");
}
__device__ __forceinline__ void mc_st128(float *p, const float *o) {
  asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(o[0]),
               "f"(o[1]), "f"(o[2]), "f"(o[3]) : "memory");
}
__device__ __forceinline__ void mc_ldred128(const float *p, float *o) {
  asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]) : "l"(p) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(float *mc, float *local, float *mc2,
                                             uint64_t shard_off, uint64_t n_elems) {
  if constexpr (MODE == 5) {
    const uint64_t n8 = n_elems / 8;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
      float v[8];
      ld256(local + i * 8, v);
      st256(mc + shard_off + i * 8, v);
    }
  } else if constexpr (MODE == 6) {
    // per warp: a 4 KiB tile (1024 floats) = 4 x 256-bit loads per lane, two tiles in flight
    extern __shared__ __align__(128) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *tile0 = smem + warp * 2048;
    const uint64_t n_tiles = n_elems / 1024;
    const uint64_t warps = static_cast<uint64_t>(gridDim.x) * (blockDim.x >> 5);
    int buf = 0;
    for (uint64_t t = static_cast<uint64_t>(blockIdx.x) * (blockDim.x >> 5) + warp; t < n_tiles;
         t += warps, buf ^= 1) {
      float *tile = tile0 + buf * 1024;
      // the bulk copy that last read this buffer (two iterations ago) must have finished reading
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncwarp();
      float v[4][8];
#pragma unroll
      for (int k = 0; k < 4; ++k) ld256(local + t * 1024 + (k * 32 + lane) * 8, v[k]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float4 *d = reinterpret_cast<float4 *>(tile + (k * 32 + lane) * 8);
        d[0] = make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
        d[1] = make_float4(v[k][4], v[k][5], v[k][6], v[k][7]);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(tile));
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                         mc + shard_off + t * 1024), "r"(s), "r"(4096) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  } else if constexpr (MODE == 8) {
    const uint64_t n4 = n_elems / 4;
    const uint64_t half = gridDim.x / 2;
    const bool reducer = (blockIdx.x & 1) == 0;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x >> 1) * blockDim.x + threadIdx.x; i < n4;
         i += half * blockDim.x) {
      float v[4];
      if (reducer) {
        mc_ldred128(mc + shard_off + i * 4, v);
        st128(local + i * 4, v);
      } else {
        ld128(local + n4 * 4 + i * 4, v);
        mc_st128(mc2 + shard_off + i * 4, v);
      }
    }
  } else {
    const uint64_t n4 = n_elems / 4;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
      float *p = mc + shard_off + i * 4;
      float v[4];
      if constexpr (MODE == 0) {
        ld128(local + i * 4, v);
        mc_st128(p, v);
      } else if constexpr (MODE == 1) {
        mc_ldred128(p, v);
        st128(local + i * 4, v);
      } else if constexpr (MODE == 2) {
        ld128(local + i * 4, v);
        st128(local + n4 * 4 + i * 4, v);
      } else if constexpr (MODE == 3) {
        ld128(p, v);
        st128(local + i * 4, v);
      } else if constexpr (MODE == 4) {
        ld128(local + i * 4, v);
        st128(p, v);
      } else {   // 7
        mc_ldred128(p, v);
        mc_st128(mc2 + shard_off + i * 4, v);
      }
    }
  }
}

}  // namespace

extern "C" int accel_probe_run(float *mc, float *local, float *mc2, uint64_t shard_off_elems,
                               uint64_t n_elems, int mode, int ctas, void *stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (mode) {
    case 0: probe<0><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 1: probe<1><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 2: probe<2><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 3: probe<3><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 4: probe<4><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 5: probe<5><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 6: {
      static bool attr = false;
      if (!attr) {
        cudaFuncSetAttribute(probe<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
        attr = true;
      }
      probe<6><<<ctas, 256, 65536, s>>>(mc, local, mc2, shard_off_elems, n_elems);
      break;
    }
    case 7: probe<7><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    case 8: probe<8><<<ctas, 256, 0, s>>>(mc, local, mc2, shard_off_elems, n_elems); break;
    default: return -1;
  }
  return cudaGetLastError() == cudaSuccess ? 0 : -3;
}
