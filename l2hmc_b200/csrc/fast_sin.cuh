// Branch-free fp32 sin / sincos for the fused HMC kernels.
//
// Plaquette angles are sums of four links and stay within a few tens of radians, so the full
// libdevice sinf (Payne-Hanek slow path, quadrant selects: ~40 SASS instructions per call, the
// majority of the fused kernels' issue slots in the r1 ncu capture) is replaced by:
//   j = rint(x / 2pi)            (magic-number add, FMA pipe)
//   r = x - j*2pi                (3-term Cody-Waite, 2pi = C1 + C2 + C3, exact products for |j| < 2^13)
//   fold |r| > pi/2 to pi - |r|  (pi = hi + lo)
//   odd / even minimax polynomial on [-pi/2, pi/2]
// ~17 instructions.  Max ABSOLUTE error 1.6e-7 (sin) / 1.9e-7 (cos) for |x| <= 4e4, validated
// against float64 in scripts/design_sin.py (coefficients generated there); |x| > 4e4 falls back
// to sinf/cosf.  The force needs absolute accuracy (beta * sin P), which this matches to
// ~1.3 ulp of 1.0 -- the same size as the difference between two correctly-rounded libm's.
#pragma once

namespace l2hmc {

struct SinReduce {
  float r;     // folded argument in [-pi/2, pi/2] (sign of the reduced argument applied)
  float a;     // |folded|
  bool fold;   // pi - |r| was taken (cos changes sign)
};

__device__ __forceinline__ SinReduce sin_reduce(float x) {
  const float magic = 12582912.0f;
  const float j = __fsub_rn(fmaf(x, 0.15915494309189535f, magic), magic);
  float r = fmaf(j, -6.28125f, x);
  r = fmaf(j, -0.0019350051879882812f, r);
  r = fmaf(j, -3.019916050561733e-07f, r);
  const float a = fabsf(r);
  const float folded = __fadd_rn(__fsub_rn(3.14159274101257324f, a), -8.742277657347586e-08f);
  SinReduce o;
  o.fold = a > 1.57079637050628662f;
  o.a = o.fold ? folded : a;
  // copy the sign of r onto the (possibly slightly negative) folded value: xor, not copysign
  o.r = __int_as_float(__float_as_int(o.a) ^ (__float_as_int(r) & 0x80000000));
  return o;
}

__device__ __forceinline__ float sin_poly(float r) {
  const float s = __fmul_rn(r, r);
  float p = 2.599861772978329e-06f;
  p = fmaf(p, s, -0.0001980652887141332f);
  p = fmaf(p, s, 0.008333016186952591f);
  p = fmaf(p, s, -0.16666656732559204f);
  return fmaf(__fmul_rn(r, s), p, r);
}

__device__ __forceinline__ float cos_poly(float a) {
  const float s = __fmul_rn(a, a);
  float p = -2.607527846976154e-07f;
  p = fmaf(p, s, 2.476178815413732e-05f);
  p = fmaf(p, s, -0.0013888401445001364f);
  p = fmaf(p, s, 0.04166664183139801f);
  p = fmaf(p, s, -0.5f);
  return fmaf(s, p, 1.0f);
}

__device__ __forceinline__ float fast_sin(float x) {
  if (__builtin_expect(fabsf(x) > 40000.f, 0)) return sinf(x);
  return sin_poly(sin_reduce(x).r);
}

__device__ __forceinline__ float fast_cos(float x) {
  if (__builtin_expect(fabsf(x) > 40000.f, 0)) return cosf(x);
  const SinReduce q = sin_reduce(x);
  const float c = cos_poly(q.a);
  return q.fold ? -c : c;
}

__device__ __forceinline__ void fast_sincos(float x, float* sn, float* cs) {
  if (__builtin_expect(fabsf(x) > 40000.f, 0)) {
    sincosf(x, sn, cs);
    return;
  }
  const SinReduce q = sin_reduce(x);
  *sn = sin_poly(q.r);
  const float c = cos_poly(q.a);
  *cs = q.fold ? -c : c;
}

// sin, cos and the projection of P onto [-pi, pi) from ONE range reduction (the trajectory ends of the fused MC-step
// kernel need all three per plaquette).  sin and cos are fast_sincos's operations; the quotient j = rint(P / 2pi) of
// the reduction is floor((P + pi) / 2pi) whenever |r| = |P - 2pi j| is clear of pi, so proj = P - 2pi j formed with the
// reference's own two roundings is bit-identical to project_angle(P) (which needs an exact division); near the branch
// cut (|r| >= 3.1415, probability ~3e-5) and for |P| > 4e4 the exact path is taken.
__device__ __forceinline__ void sincos_and_project(float P, float* sn, float* cs, float* proj) {
  if (__builtin_expect(fabsf(P) > 40000.f, 0)) {
    sincosf(P, sn, cs);
    *proj = project_angle(P);
    return;
  }
  const float magic = 12582912.0f;
  const float j = __fsub_rn(fmaf(P, 0.15915494309189535f, magic), magic);
  float r = fmaf(j, -6.28125f, P);
  r = fmaf(j, -0.0019350051879882812f, r);
  r = fmaf(j, -3.019916050561733e-07f, r);
  const float a = fabsf(r);
  const float folded = __fadd_rn(__fsub_rn(3.14159274101257324f, a), -8.742277657347586e-08f);
  const bool fold = a > 1.57079637050628662f;
  const float aa = fold ? folded : a;
  const float rr = __int_as_float(__float_as_int(aa) ^ (__float_as_int(r) & 0x80000000));
  *sn = sin_poly(rr);
  const float c = cos_poly(aa);
  *cs = fold ? -c : c;
  *proj = a < 3.1415f ? __fsub_rn(P, __fmul_rn(kTwoPi, j)) : project_angle(P);
}

// Two sines per instruction stream (FFMA2/FADD2/FMUL2): the same reduction, fold and polynomial
// as fast_sin on each lane.  The caller guarantees |x| <= 4e4 on both lanes.
__device__ __forceinline__ f32x2 fast_sin2(f32x2 x) {
  const f32x2 magic = pack2(12582912.0f, 12582912.0f);
  const f32x2 j = sub2(fma2(x, pack2(0.15915494309189535f, 0.15915494309189535f), magic), magic);
  f32x2 r = fma2(j, pack2(-6.28125f, -6.28125f), x);
  r = fma2(j, pack2(-0.0019350051879882812f, -0.0019350051879882812f), r);
  r = fma2(j, pack2(-3.019916050561733e-07f, -3.019916050561733e-07f), r);
  const f32x2 a = r & 0x7fffffff7fffffffULL;
  const f32x2 folded = add2(sub2(pack2(3.14159274101257324f, 3.14159274101257324f), a),
                            pack2(-8.742277657347586e-08f, -8.742277657347586e-08f));
  // |r| <= pi (+ rounding): min(|r|, pi - |r|) takes the fold exactly when |r| > pi/2
  const f32x2 m = pack2(fminf(lo2(a), lo2(folded)), fminf(hi2(a), hi2(folded)));
  const f32x2 rr = m ^ (r & 0x8000000080000000ULL);
  const f32x2 s = mul2(rr, rr);
  f32x2 p = fma2(pack2(2.599861772978329e-06f, 2.599861772978329e-06f), s,
                 pack2(-0.0001980652887141332f, -0.0001980652887141332f));
  p = fma2(p, s, pack2(0.008333016186952591f, 0.008333016186952591f));
  p = fma2(p, s, pack2(-0.16666656732559204f, -0.16666656732559204f));
  return fma2(mul2(rr, s), p, rr);
}

}  // namespace l2hmc
