// Per-element arithmetic of the PS step and the memory-access helpers, shared by
// every kernel variant. The arithmetic restates TF-1.4's CPU functors
// (ApplyGradientDescent / ApplyMomentum / ApplyAdamNonCuda in
// tensorflow/core/kernels/training_ops.cc — not in /root/reference; call sites
// examples/workdir/mnist_replica.py:170, mnist_softmax.py:59).
//
// Rounding: TF-1.4's stock CPU wheel evaluates these Eigen expressions with
// separately rounded SSE mul/add/div/sqrt (no FMA contraction). Every operation
// here is therefore an explicit round-to-nearest intrinsic (__fadd_rn etc. are
// never fused by nvcc), and sqrt/div are the IEEE-correct forms (the library is
// built without --use_fast_math), so fp32 results are bit-identical to the CPU
// statement-per-pass evaluation.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "ps_common.h"

namespace accel {

enum { OPT_SGD = ACCEL_OPT_SGD, OPT_MOMENTUM = ACCEL_OPT_MOMENTUM, OPT_ADAM = ACCEL_OPT_ADAM };

struct bf16_t {
  uint16_t bits;
};

// Scalars every thread needs, derived once per thread from the argument block.
struct Scalars {
  float lr, momentum, beta1, omb1, omb2, alpha, eps, mean_div, gscale;
  bool nesterov, has_div, has_scale;
};

__device__ __forceinline__ Scalars make_scalars(const ApplyArgs &a) {
  Scalars s;
  s.lr = a.lr;
  s.momentum = a.momentum;
  s.beta1 = a.beta1;
  s.omb1 = a.one_minus_beta1;
  s.omb2 = a.one_minus_beta2;
  s.alpha = a.alpha;
  if (a.dev_powers != nullptr) {
    // beta powers live on the device: alpha = lr * sqrt(1 - beta2^t) / (1 - beta1^t), the same
    // fp32 operations in the same order as the host computes it
    const float b1p = __ldg(a.dev_powers), b2p = __ldg(a.dev_powers + 1);
    s.alpha = __fdiv_rn(__fmul_rn(a.lr, __fsqrt_rn(__fsub_rn(1.0f, b2p))), __fsub_rn(1.0f, b1p));
  }
  s.eps = a.epsilon;
  s.mean_div = a.mean_div;
  s.nesterov = a.use_nesterov != 0;
  s.has_div = a.mean_div != 0.0f;
  s.has_scale = false;
  s.gscale = 1.0f;
  if (a.clip_norm > 0.0f && a.sumsq != nullptr) {
    // tf.clip_by_global_norm (TF-1.4 clip_ops.py): scale = clip_norm *
    // min(1/global_norm, 1/clip_norm), global_norm = sqrt(sum g^2).
    float gn = __fsqrt_rn(__ldg(a.sumsq));
    float inv_gn = __fdiv_rn(1.0f, gn);
    float inv_c = __fdiv_rn(1.0f, a.clip_norm);
    s.gscale = __fmul_rn(a.clip_norm, fminf(inv_gn, inv_c));
    s.has_scale = true;
  }
  return s;
}

// The summed gradient -> the gradient the optimizer sees.
__device__ __forceinline__ float finish_grad(float sum, const Scalars &s) {
  if (s.has_div) sum = __fdiv_rn(sum, s.mean_div);      // accumulator take_grad: acc / N
  if (s.has_scale) sum = __fmul_rn(sum, s.gscale);      // global-norm clip
  return sum;
}

template <int OPT>
__device__ __forceinline__ void update(float &var, float &s0, float &s1, float g,
                                       const Scalars &s) {
  if constexpr (OPT == OPT_SGD) {
    // var -= grad * lr
    var = __fsub_rn(var, __fmul_rn(g, s.lr));
  } else if constexpr (OPT == OPT_MOMENTUM) {
    // accum = accum * momentum + grad
    s0 = __fadd_rn(__fmul_rn(s0, s.momentum), g);
    if (s.nesterov) {
      // var -= grad * lr + accum * momentum * lr
      float t = __fadd_rn(__fmul_rn(g, s.lr), __fmul_rn(__fmul_rn(s0, s.momentum), s.lr));
      var = __fsub_rn(var, t);
    } else {
      // var -= accum * lr
      var = __fsub_rn(var, __fmul_rn(s0, s.lr));
    }
  } else {
    // m += (grad - m) * (1 - beta1)
    s0 = __fadd_rn(s0, __fmul_rn(__fsub_rn(g, s0), s.omb1));
    // v += (grad^2 - v) * (1 - beta2)
    s1 = __fadd_rn(s1, __fmul_rn(__fsub_rn(__fmul_rn(g, g), s1), s.omb2));
    float denom = __fadd_rn(__fsqrt_rn(s1), s.eps);
    if (s.nesterov) {
      // var -= ((grad*(1-beta1) + beta1*m) * alpha) / (sqrt(v) + eps)
      float num = __fmul_rn(__fadd_rn(__fmul_rn(g, s.omb1), __fmul_rn(s.beta1, s0)), s.alpha);
      var = __fsub_rn(var, __fdiv_rn(num, denom));
    } else {
      // var -= (m * alpha) / (sqrt(v) + eps)
      var = __fsub_rn(var, __fdiv_rn(__fmul_rn(s0, s.alpha), denom));
    }
  }
}

template <int OPT>
struct OptTraits;
template <>
struct OptTraits<OPT_SGD> {
  static constexpr int kSlots = 0;
};
template <>
struct OptTraits<OPT_MOMENTUM> {
  static constexpr int kSlots = 1;
};
template <>
struct OptTraits<OPT_ADAM> {
  static constexpr int kSlots = 2;
};

// ---- memory access --------------------------------------------------------
// Gradients are read exactly once: non-coherent path, no L1 allocation.
__device__ __forceinline__ void ldg_stream_128(const void *p, float (&o)[4]) {
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
      : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3])
      : "l"(p));
}
__device__ __forceinline__ void ldg_stream_128u(const void *p, uint32_t (&o)[4]) {
  asm("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
      : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3])
      : "l"(p));
}
__device__ __forceinline__ void ldg_stream_256(const void *p, float (&o)[8]) {
  asm("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]), "=f"(o[4]), "=f"(o[5]), "=f"(o[6]),
        "=f"(o[7])
      : "l"(p));
}
// State (var / slots) is read-modify-written by the same thread.
__device__ __forceinline__ void ld_state_128(const float *p, float *o) {
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3])
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void st_state_128(float *p, const float *o) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(o[0]),
               "f"(o[1]), "f"(o[2]), "f"(o[3])
               : "memory");
}
__device__ __forceinline__ void ld_state_256(const float *p, float *o) {
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]), "=f"(o[4]), "=f"(o[5]),
                 "=f"(o[6]), "=f"(o[7])
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void st_state_256(float *p, const float *o) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
               "f"(o[0]), "f"(o[1]), "f"(o[2]), "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]),
               "f"(o[7])
               : "memory");
}

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// Element type traits. VE = elements one thread handles per vector item.
template <typename G, bool WIDE>
struct Item;
template <>
struct Item<float, false> {
  static constexpr int VE = 4;
};
template <>
struct Item<float, true> {
  static constexpr int VE = 8;
};
template <bool WIDE>
struct Item<bf16_t, WIDE> {
  static constexpr int VE = 8;
};

// Load VE gradient elements of item `item` from one worker's buffer, as fp32.
template <typename G, bool WIDE>
__device__ __forceinline__ void load_grad_item(const void *base, uint64_t item,
                                               float (&o)[Item<G, WIDE>::VE]) {
  if constexpr (sizeof(G) == 4) {
    if constexpr (WIDE) {
      ldg_stream_256(static_cast<const float *>(base) + item * 8, o);
    } else {
      ldg_stream_128(static_cast<const float *>(base) + item * 4, o);
    }
  } else {
    uint32_t u[4];
    ldg_stream_128u(static_cast<const uint16_t *>(base) + item * 8, u);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[2 * k] = bf16_lo(u[k]);       // bf16 -> fp32 is exact
      o[2 * k + 1] = bf16_hi(u[k]);
    }
  }
}

// ---- the apply kernel's view: gradients stay PACKED in registers until they are summed -----
// VE = elements per item, R = 32-bit registers one worker's item occupies. Wide items are one
// 256-bit load per worker: 8 fp32 or 16 bf16 elements.
template <typename G, bool WIDE, int OPT>
struct ApplyItem;
template <bool WIDE, int OPT>
struct ApplyItem<float, WIDE, OPT> {
  static constexpr int VE = WIDE ? 8 : 4;
  static constexpr int R = VE;
};
// bf16: 16 elements (one 256-bit gradient load) measured faster for SGD and Momentum (0.370 vs
// 0.380 ms, 0.479 vs 0.493 ms at 8 x 100 M) and slower for Adam, whose three state arrays then
// need 48 registers per item (0.645 vs 0.628 ms) — Adam keeps 8-element items.
template <bool WIDE, int OPT>
struct ApplyItem<bf16_t, WIDE, OPT> {
  static constexpr int VE = (WIDE && OPT != OPT_ADAM) ? 16 : 8;
  static constexpr int R = VE / 2;
};

__device__ __forceinline__ void ldg_stream_256u(const void *p, uint32_t (&o)[8]) {
  asm("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3]), "=r"(o[4]), "=r"(o[5]), "=r"(o[6]),
        "=r"(o[7])
      : "l"(p));
}

template <int R>
__device__ __forceinline__ void load_grad_raw(const void *base, uint64_t item, uint32_t (&o)[R]) {
  const uint32_t *p = static_cast<const uint32_t *>(base) + item * R;
  if constexpr (R == 8) {
    ldg_stream_256u(p, o);
  } else {
    ldg_stream_128u(p, o);
  }
}

// element e of a packed item as fp32 (bf16 -> fp32 is exact)
template <typename G>
__device__ __forceinline__ float grad_elem(const uint32_t *raw, int e) {
  if constexpr (sizeof(G) == 4) {
    return __uint_as_float(raw[e]);
  } else {
    return (e & 1) ? bf16_hi(raw[e >> 1]) : bf16_lo(raw[e >> 1]);
  }
}

template <typename G>
__device__ __forceinline__ float load_grad_scalar(const void *base, uint64_t i) {
  if constexpr (sizeof(G) == 4) {
    return __ldg(static_cast<const float *>(base) + i);
  } else {
    return __uint_as_float(static_cast<uint32_t>(__ldg(static_cast<const uint16_t *>(base) + i))
                           << 16);
  }
}

template <int VE, bool WIDE>
__device__ __forceinline__ void load_state_item(const float *base, uint64_t item, float (&o)[VE]) {
  const float *p = base + item * VE;
  if constexpr (VE == 16) {
    if constexpr (WIDE) {
      ld_state_256(p, o);
      ld_state_256(p + 8, o + 8);
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) ld_state_128(p + 4 * k, o + 4 * k);
    }
  } else if constexpr (VE == 8 && WIDE) {
    ld_state_256(p, o);
  } else if constexpr (VE == 8) {
    ld_state_128(p, o);
    ld_state_128(p + 4, o + 4);
  } else {
    ld_state_128(p, o);
  }
}
template <int VE, bool WIDE>
__device__ __forceinline__ void store_state_item(float *base, uint64_t item,
                                                 const float (&o)[VE]) {
  float *p = base + item * VE;
  if constexpr (VE == 16) {
    if constexpr (WIDE) {
      st_state_256(p, o);
      st_state_256(p + 8, o + 8);
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) st_state_128(p + 4 * k, o + 4 * k);
    }
  } else if constexpr (VE == 8 && WIDE) {
    st_state_256(p, o);
  } else if constexpr (VE == 8) {
    st_state_128(p, o);
    st_state_128(p + 4, o + 4);
  } else {
    st_state_128(p, o);
  }
}

}  // namespace accel
