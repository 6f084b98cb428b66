// Per-element arithmetic of the four augmented-leapfrog sub-updates, shared by the stand-alone
// K2 kernels (updates.cu) and the network head epilogues (nets.cu, gemm_tc.cu).
// Every product and sum is rounded separately in the reference's association order
// (reference l2hmc/dynamics/gauge_dynamics.py:486-590); exp() is expf (<= 1 ulp).
#pragma once
#include "common.cuh"

namespace l2hmc {

// dir = +1: v' = v*exp(eps/2*S) - eps/2*(exp(eps*Q)*g - T)        (:486-508)
// dir = -1: v' = exp(-eps/2*S)*(v + eps/2*(exp(eps*Q)*g - T))     (:537-561)
// logdet += (+-)eps/2*S
template <int DIR>
__device__ __forceinline__ float momentum_elem(float v, float g, float S, float Tr, float Q, float eps, float heps,
                                               float& logdet, bool& bad) {
  const float sc = __fmul_rn(S, DIR > 0 ? heps : -heps);
  const float se = expf(sc);
  const float te = expf(__fmul_rn(Q, eps));
  bad |= !(isfinite(se) && isfinite(te));
  logdet = __fadd_rn(logdet, sc);
  const float inner = __fsub_rn(__fmul_rn(te, g), Tr);
  if (DIR > 0) return __fsub_rn(__fmul_rn(v, se), __fmul_rn(heps, inner));
  return __fmul_rn(se, __fadd_rn(v, __fmul_rn(heps, inner)));
}

// dir = +1: x' = m*x + mbar*(x*exp(eps*S) + eps*(exp(eps*Q)*v + T))      (:511-534)
// dir = -1: x' = m*x + mbar*exp(-eps*S)*(x - eps*(exp(eps*Q)*v + T))     (:565-590)
// logdet += mbar*(+-)eps*S
template <int DIR>
__device__ __forceinline__ float position_elem(float x, float v, float S, float Tr, float Q, float m, float eps,
                                               float& logdet, bool& bad) {
  const float mi = __fsub_rn(1.f, m);
  const float sc = __fmul_rn(S, DIR > 0 ? eps : -eps);
  const float se = expf(sc);
  const float te = expf(__fmul_rn(Q, eps));
  bad |= !(isfinite(se) && isfinite(te));
  logdet = __fadd_rn(logdet, __fmul_rn(mi, sc));
  const float inner = __fadd_rn(__fmul_rn(te, v), Tr);
  float tmp;
  if (DIR > 0) tmp = __fadd_rn(__fmul_rn(x, se), __fmul_rn(eps, inner));
  else tmp = __fmul_rn(se, __fsub_rn(x, __fmul_rn(eps, inner)));
  return __fadd_rn(__fmul_rn(m, x), __fmul_rn(mi, tmp));
}

// warp-uniform: OR 1 into the device flag when any lane saw a non-finite exp()
// (device-side mirror of tf.check_numerics, gauge_dynamics.py:26-28)
__device__ __forceinline__ void flag_nonfinite(int32_t* flag, bool bad) {
  if (flag != nullptr && __any_sync(0xffffffffu, bad)) {
    if ((threadIdx.x & 31) == 0) atomicOr(flag, 1);
  }
}

}  // namespace l2hmc
