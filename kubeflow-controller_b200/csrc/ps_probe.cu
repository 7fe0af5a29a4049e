// Development probes (not part of the public ABI): raw rates of the NVSwitch multicast
// operations a fused kernel could use. tools/nvls_probe.py drives them under torchrun.
#include "ps_shard.cuh"

namespace accel {

// mode 0: multimem.st of this rank's shard through the multicast mapping (all-gather pattern)
// mode 1: multimem.ld_reduce of this rank's shard (in-switch sum over all ranks), result to `local`
// mode 2: plain local copy of the shard (HBM reference)
// mode 3 / 4: unicast peer load / peer store of the same bytes (`mc` = the peer's buffer)
template <int MODE>
__global__ void __launch_bounds__(256) ps_probe_multimem(float *mc, float *local, uint64_t shard_off,
                                                         uint64_t n_vec4) {
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_vec4;
       i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    float *p = mc + shard_off + i * 4;
    float v[4];
    if constexpr (MODE == 0) {
      ld_state_128(local + i * 4, v);
      asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v[0]),
                   "f"(v[1]), "f"(v[2]), "f"(v[3])
                   : "memory");
    } else if constexpr (MODE == 1) {
      asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3])
                   : "l"(p)
                   : "memory");
      st_state_128(local + i * 4, v);
    } else if constexpr (MODE == 2) {
      ld_state_128(local + i * 4, v);
      st_state_128(local + n_vec4 * 4 + i * 4, v);
    } else if constexpr (MODE == 3) {
      // unicast peer LOAD: `mc` is a peer's buffer here; read its shard, store locally
      ld_state_128(p, v);
      st_state_128(local + i * 4, v);
    } else {
      // unicast peer STORE: read locally, store into the peer's buffer
      ld_state_128(local + i * 4, v);
      st_state_128(p, v);
    }
  }
}

}  // namespace accel

extern "C" int accel_dev_probe_multimem(float *mc, float *local, uint64_t shard_off_elems,
                                        uint64_t n_elems, int mode, int ctas, void *stream) {
  using namespace accel;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint64_t n4 = n_elems / 4;
  switch (mode) {
    case 0: ps_probe_multimem<0><<<ctas, 256, 0, s>>>(mc, local, shard_off_elems, n4); break;
    case 1: ps_probe_multimem<1><<<ctas, 256, 0, s>>>(mc, local, shard_off_elems, n4); break;
    case 3: ps_probe_multimem<3><<<ctas, 256, 0, s>>>(mc, local, shard_off_elems, n4); break;
    case 4: ps_probe_multimem<4><<<ctas, 256, 0, s>>>(mc, local, shard_off_elems, n4); break;
    default: ps_probe_multimem<2><<<ctas, 256, 0, s>>>(mc, local, shard_off_elems, n4); break;
  }
  ACCEL_CUDA(cudaGetLastError());
  return ACCEL_OK;
}
