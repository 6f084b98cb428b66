// K3 conv front-end of ConvNet3D: per input stream
//   (N,T,X,2,1) -> Conv3D(F, (3,3,2), 'same', relu) -> MaxPool3D(2, 2, 'same')
//               -> Conv3D(2F, (2,2,2), 'same', relu) -> MaxPool3D(2, 2, 'same') -> flatten
// in ONE kernel, one chain per CTA pass with everything on-chip.
//
// Replaces reference l2hmc/network/conv_net.py:247-260 (+ reshape_5D :283-311, layers :86-168),
// data_format = channels_last.  TensorFlow 'same' semantics: ZERO padding (not periodic), total
// pad = k - 1 with the odd cell at the END (so the 2-wide kernels only see pad after), and the
// pooling windows are clipped at the end (-inf padding).  The lattice's link direction mu is the
// third spatial axis (size 2); after the first pool it has size 1, so the kd = 1 taps of the
// second convolution only ever see padding and are skipped.
//
// Work split: phase 1 computes pooled conv-1 outputs P1[t1][x1][f] (f fastest: weight reads are
// conflict-free, input reads are broadcasts) into shared memory; phase 2 computes the four
// conv-2 positions of a pooling window together from a 3x3 patch of P1 held in registers.
// relu(max(.)) == max(relu(.)), and max(a_i + b) == max(a_i) + b in floating point (monotone),
// so bias and relu are applied once per pooled output.
#include <algorithm>
#include <cstdlib>
#include <string>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace l2hmc {
namespace {

constexpr int kThreads = 256;

struct ConvParams {
  const float* w1;   // [3][3][2][1][F]
  const float* b1;   // [F]
  const float* w2;   // [2][2][2][F][2F]
  const float* b2;   // [2F]
  const float* in;   // [N][T][X][2]
  const float* mask; // nullable [T*X*2]
  int invert;
  const float* mask_alt;     // chains with row_set[n] != 0 use this pair instead
  int invert_alt;
  const uint8_t* row_set;    // nullable [N]
  int N, T, X, F;
  float* feat;       // [N][T2*X2*2F]
  // routing tape for the training pass (l2hmc_conv3d_tape), all three nullable together
  float* t_p1;       // [N][T1*X1*F]
  uint8_t* t_a1;     // [N][T1*X1*F]
  uint8_t* t_a2;     // [N][T2*X2*2F]
};

__global__ void __launch_bounds__(kThreads) conv3d_kernel(const ConvParams p) {
  extern __shared__ __align__(16) float sm[];
  const int T = p.T, X = p.X, F = p.F, C = 2 * p.F;
  const int T1 = (T + 1) / 2, X1 = (X + 1) / 2, T2 = (T1 + 1) / 2, X2 = (X1 + 1) / 2;
  const int D = 2 * T * X;
  float* s_w1 = sm;                       // [18][F]
  float* s_b1 = s_w1 + 18 * F;            // [F]
  float* s_w2 = s_b1 + F;                 // [4][F][C]   (kd = 0 taps only)
  float* s_b2 = s_w2 + 4 * F * C;         // [C]
  float* s_in = sm + ((18 * F + F + 4 * F * C + C + 3) & ~3);   // [T][X][2]  (float2 loads: aligned also for odd F)
  float* s_p1 = s_in + ((D + 3) & ~3);    // [T1][X1][F]
  const int tid = threadIdx.x;

  for (int i = tid; i < 18 * F; i += kThreads) s_w1[i] = p.w1[i];
  for (int i = tid; i < F; i += kThreads) s_b1[i] = p.b1[i];
  for (int i = tid; i < 4 * F * C; i += kThreads) {
    const int tap = i / (F * C), rest = i % (F * C);        // tap = kt*2 + kx; source index [kt][kx][kd=0][f][c]
    s_w2[i] = p.w2[static_cast<size_t>(tap) * 2 * F * C + rest];
  }
  for (int i = tid; i < C; i += kThreads) s_b2[i] = p.b2[i];

  for (int n = blockIdx.x; n < p.N; n += gridDim.x) {
    __syncthreads();   // previous chain's phase 2 done with s_p1 / s_in; weights visible on the first pass
    const float* src = p.in + static_cast<size_t>(n) * D;
    const bool alt = p.row_set != nullptr && p.row_set[n] != 0;
    const float* mask = alt ? p.mask_alt : p.mask;
    const int invert = alt ? p.invert_alt : p.invert;
    for (int i = tid; i < D; i += kThreads) {
      float a = src[i];
      if (mask) {
        const float m = mask[i];
        a = __fmul_rn(invert ? __fsub_rn(1.f, m) : m, a);
      }
      s_in[i] = a;
    }
    __syncthreads();

    // ---- phase 1: conv (3,3,2) + bias + relu + pool (2,2,2)
    for (int o = tid; o < T1 * X1 * F; o += kThreads) {
      const int f = o % F, pos = o / F;
      const int x1 = pos % X1, t1 = pos / X1;
      // input patch rows 2*t1-1 .. 2*t1+2, cols 2*x1-1 .. 2*x1+2, both mu (zero outside the lattice)
      float in[4][4][2];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int t = 2 * t1 - 1 + a, x = 2 * x1 - 1 + b;
          const bool ok = t >= 0 && t < T && x >= 0 && x < X;
          const float2 u = ok ? *reinterpret_cast<const float2*>(&s_in[(t * X + x) * 2]) : make_float2(0.f, 0.f);
          in[a][b][0] = u.x, in[a][b][1] = u.y;
        }
      float w[3][3][2];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
          w[a][b][0] = s_w1[((a * 3 + b) * 2 + 0) * F + f];
          w[a][b][1] = s_w1[((a * 3 + b) * 2 + 1) * F + f];
        }
      float best = -INFINITY;
      int arg = 0;                                                 // first maximum in (dt, dx, mu) order
#pragma unroll
      for (int dt = 0; dt < 2; ++dt)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx) {
          if (2 * t1 + dt >= T || 2 * x1 + dx >= X) continue;     // pooling window clipped at the end
          float s0 = 0.f, s1 = 0.f;                                // outputs at mu = 0 and mu = 1
#pragma unroll
          for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) {
              // mu = 0 sees taps kd = 0 (mu 0) and kd = 1 (mu 1); mu = 1 sees kd = 0 (mu 1), kd = 1 -> pad
              s0 = fmaf(in[dt + a][dx + b][0], w[a][b][0], s0);
              s0 = fmaf(in[dt + a][dx + b][1], w[a][b][1], s0);
              s1 = fmaf(in[dt + a][dx + b][1], w[a][b][0], s1);
            }
          if (s0 > best) best = s0, arg = dt * 4 + dx * 2;
          if (s1 > best) best = s1, arg = dt * 4 + dx * 2 + 1;
        }
      const float pooled = fmaxf(__fadd_rn(best, s_b1[f]), 0.f);
      s_p1[o] = pooled;
      if (p.t_p1 != nullptr) {
        p.t_p1[static_cast<size_t>(n) * T1 * X1 * F + o] = pooled;
        p.t_a1[static_cast<size_t>(n) * T1 * X1 * F + o] = pooled > 0.f ? static_cast<uint8_t>(arg) : 255;
      }
    }
    __syncthreads();

    // ---- phase 2: conv (2,2,[2]) + bias + relu + pool (2,2,[1]) -> flattened features
    float* dst = p.feat + static_cast<size_t>(n) * T2 * X2 * C;
    for (int o = tid; o < T2 * X2 * C; o += kThreads) {
      const int c = o % C, pos = o / C;
      const int x2 = pos % X2, t2 = pos / X2;
      float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
      const int tb = 2 * t2, xb = 2 * x2;
      for (int f = 0; f < F; ++f) {
        float pp[3][3];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int b = 0; b < 3; ++b) {
            const int t = tb + a, x = xb + b;
            pp[a][b] = (t < T1 && x < X1) ? s_p1[(t * X1 + x) * F + f] : 0.f;
          }
        const float w00 = s_w2[(0 * F + f) * C + c], w01 = s_w2[(1 * F + f) * C + c];
        const float w10 = s_w2[(2 * F + f) * C + c], w11 = s_w2[(3 * F + f) * C + c];
#pragma unroll
        for (int dt = 0; dt < 2; ++dt)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            float s = acc[dt][dx];
            s = fmaf(pp[dt][dx], w00, s);
            s = fmaf(pp[dt][dx + 1], w01, s);
            s = fmaf(pp[dt + 1][dx], w10, s);
            s = fmaf(pp[dt + 1][dx + 1], w11, s);
            acc[dt][dx] = s;
          }
      }
      float best = -INFINITY;
      int arg = 0;
#pragma unroll
      for (int dt = 0; dt < 2; ++dt)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx)
          if (tb + dt < T1 && xb + dx < X1 && acc[dt][dx] > best) best = acc[dt][dx], arg = dt * 2 + dx;
      const float pooled = fmaxf(__fadd_rn(best, s_b2[c]), 0.f);
      dst[o] = pooled;
      if (p.t_a2 != nullptr) p.t_a2[static_cast<size_t>(n) * T2 * X2 * C + o] = pooled > 0.f ? static_cast<uint8_t>(arg) : 255;
    }
  }
}


// ---------------------------------------------------------------------------------------------
// Register-tiled, packed-fp32x2 variant for F in {8, 16, 32} and T, X multiples of 4 (BASELINE configs
// 2-4 have F = X = T): kChains chains per CTA pass, no boundary logic in the arithmetic, and every
// multiply-add is one lane of an FFMA2 (two FMAs per thread per instruction: the plain FFMA pipe tops
// out at half the FP32 rate on sm_100).
//   staging: the NEXT pass's links are fetched into registers (PF float4 per thread) before this
//            pass computes, and written -- mask applied, each value DUPLICATED (a, a) so a 64-bit load is a
//            ready broadcast operand -- into a zero-bordered (T+2) x (X+2) plane: TF's 'same' zero
//            padding is a plain load.
//   phase 1: a thread keeps the 18 taps of TWO neighbouring filters (f, f+1) as pairs in registers for
//            the whole kernel and streams the 4x4 input patch of one pooled output row by row
//            (16 LDS.128 per 108 FFMA2); the pooled value goes to P1, duplicated.
//   phase 2: a thread owns one pooled output position x four output channels (two pairs): per pair of
//            input filters it reads the 3x3 P1 patch (9 LDS.128) and the weights as channel pairs
//            (8 LDS.128) for 64 FFMA2.  P1 has a zero row / column at the end (the 2-wide kernel's padding).
// Per output the taps are accumulated in the generic kernel's order, so the results are bit-identical to it.
template <int F, int PF, int CH, int TH>
__global__ void __launch_bounds__(TH, 512 / TH) conv3d_tiled_kernel(const ConvParams p) {
  extern __shared__ __align__(16) float sm[];
  constexpr int C = 2 * F, FPD = 2 * F + 8, CG = C / 4, FH = F / 2, G = TH / FH;
  const int T = p.T, X = p.X;
  const int T1 = T / 2, X1 = X / 2, T2 = T1 / 2, X2 = X1 / 2;
  const int D = 2 * T * X, XB = X + 2, IN = (T + 2) * XB * 4;          // bordered plane: [t][x][mu][dup]
  const int X1B = X1 + 1, P1 = (T1 + 1) * X1B * FPD;                  // P1: [t1][x1][f][dup] + zero end row / column
  float* s_w2 = sm;                        // [4][F][C]   (kd = 0 taps only)
  float* s_b2 = s_w2 + 4 * F * C;          // [C]
  float* s_in = s_b2 + C;                  // [CH][IN]
  float* s_p1 = s_in + CH * IN;       // [CH][P1]
  const int tid = threadIdx.x;

  for (int i = tid; i < 4 * F * C; i += TH) {
    const int tap = i / (F * C), rest = i % (F * C);
    s_w2[i] = p.w2[static_cast<size_t>(tap) * 2 * F * C + rest];
  }
  for (int i = tid; i < C; i += TH) s_b2[i] = p.b2[i];
  for (int i = tid; i < CH * IN; i += TH) s_in[i] = 0.f;    // borders stay zero for the whole kernel
  for (int i = tid; i < CH * P1; i += TH) s_p1[i] = 0.f;
  const int fp = tid % FH, grp = tid / FH;                             // filters 2 fp, 2 fp + 1
  f32x2 w[3][3][2];
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < 3; ++b)
#pragma unroll
      for (int m = 0; m < 2; ++m)
        w[a][b][m] = pack2(p.w1[((a * 3 + b) * 2 + m) * F + 2 * fp], p.w1[((a * 3 + b) * 2 + m) * F + 2 * fp + 1]);
  const float bias1a = p.b1[2 * fp], bias1b = p.b1[2 * fp + 1];
  const int quads = D / 4;                 // float4 per chain

  float4 pre[PF];
  auto fetch = [&](int n0) {               // links of chains n0 .. n0 + CH - 1 -> registers
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i4 = tid + q * TH;
      const int ch = i4 / quads;
      pre[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ch < CH && n0 + ch < p.N)
        pre[q] = __ldcs(reinterpret_cast<const float4*>(p.in + static_cast<size_t>(n0 + ch) * D) + (i4 - ch * quads));
    }
  };
  auto stage = [&](int n0) {               // registers -> bordered plane, mask applied, values duplicated
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i4 = tid + q * TH;
      const int ch = i4 / quads;
      if (ch >= CH || n0 + ch >= p.N) continue;
      const int j = (i4 - ch * quads) * 4;
      float4 a = pre[q];
      const bool alt = p.row_set != nullptr && p.row_set[n0 + ch] != 0;
      const float* mask = alt ? p.mask_alt : p.mask;
      if (mask) {
        float4 m = *reinterpret_cast<const float4*>(mask + j);
        if (alt ? p.invert_alt : p.invert)
          m = make_float4(__fsub_rn(1.f, m.x), __fsub_rn(1.f, m.y), __fsub_rn(1.f, m.z), __fsub_rn(1.f, m.w));
        a = make_float4(__fmul_rn(m.x, a.x), __fmul_rn(m.y, a.y), __fmul_rn(m.z, a.z), __fmul_rn(m.w, a.w));
      }
      const int site = j >> 1, t = site / X, x = site - t * X;       // two sites (x, x + 1) of one row
      float* dst = s_in + ch * IN + ((t + 1) * XB + x + 1) * 4;
      *reinterpret_cast<float4*>(dst) = make_float4(a.x, a.x, a.y, a.y);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(a.z, a.z, a.w, a.w);
    }
  };

  int n0 = blockIdx.x * CH;
  if (n0 < p.N) fetch(n0);
  for (; n0 < p.N; n0 += gridDim.x * CH) {
    const int nch = min(CH, p.N - n0);
    __syncthreads();                       // previous pass done with s_in / s_p1 (and the zero fill on the first)
    stage(n0);
    const int n_next = n0 + gridDim.x * CH;
    if (n_next < p.N) fetch(n_next);       // in flight while this pass computes
    __syncthreads();

    // ---- phase 1: one pooled output (t1, x1) of filters (2 fp, 2 fp + 1) per item
    for (int item = grp; item < nch * T1 * X1; item += G) {
      const int x1 = item % X1, rest = item / X1, t1 = rest % T1, ch = rest / T1;
      const float* in_c = s_in + ch * IN + (2 * t1 * XB + 2 * x1) * 4;   // bordered coords of (2 t1 - 1, 2 x1 - 1)
      f32x2 s0[2][2], s1[2][2];            // conv outputs at mu = 0 / mu = 1 of the four window positions
#pragma unroll
      for (int i = 0; i < 4; ++i) (&s0[0][0])[i] = 0ull, (&s1[0][0])[i] = 0ull;
#pragma unroll
      for (int r = 0; r < 4; ++r) {        // patch row r feeds window row dt through tap row a = r - dt
        ulonglong2 row[4];                 // .x = (mu0, mu0), .y = (mu1, mu1)
#pragma unroll
        for (int b = 0; b < 4; ++b) row[b] = *reinterpret_cast<const ulonglong2*>(&in_c[(r * XB + b) * 4]);
#pragma unroll
        for (int dt = 0; dt < 2; ++dt) {
          const int a = r - dt;
          if (a < 0 || a > 2) continue;
#pragma unroll
          for (int dx = 0; dx < 2; ++dx)
#pragma unroll
            for (int b = 0; b < 3; ++b) {
              // mu = 0 sees taps kd = 0 (mu 0) and kd = 1 (mu 1); mu = 1 sees kd = 0 (mu 1), kd = 1 -> pad
              s0[dt][dx] = fma2(row[dx + b].x, w[a][b][0], s0[dt][dx]);
              s0[dt][dx] = fma2(row[dx + b].y, w[a][b][1], s0[dt][dx]);
              s1[dt][dx] = fma2(row[dx + b].y, w[a][b][0], s1[dt][dx]);
            }
        }
      }
      float ba = -INFINITY, bb = -INFINITY;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ba = fmaxf(ba, fmaxf(lo2((&s0[0][0])[i]), lo2((&s1[0][0])[i])));
        bb = fmaxf(bb, fmaxf(hi2((&s0[0][0])[i]), hi2((&s1[0][0])[i])));
      }
      ba = fmaxf(__fadd_rn(ba, bias1a), 0.f);
      bb = fmaxf(__fadd_rn(bb, bias1b), 0.f);
      *reinterpret_cast<float4*>(&s_p1[ch * P1 + (t1 * X1B + x1) * FPD + 4 * fp]) = make_float4(ba, ba, bb, bb);
    }
    __syncthreads();

    // ---- phase 2
    const int cg = tid % CG;
    for (int item = tid / CG; item < nch * T2 * X2; item += TH / CG) {
      const int x2 = item % X2, rest = item / X2, t2 = rest % T2, ch = rest / T2;
      const float* p1 = s_p1 + ch * P1 + (2 * t2 * X1B + 2 * x2) * FPD;
      f32x2 acc[2][2][2];                  // [dt][dx][channel pair]
#pragma unroll
      for (int i = 0; i < 8; ++i) (&acc[0][0][0])[i] = 0ull;
#pragma unroll 2
      for (int f2 = 0; f2 < FH; ++f2) {
        ulonglong2 pp[3][3];               // .x = (P1[f], P1[f]) of f = 2 f2, .y of f = 2 f2 + 1
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int b = 0; b < 3; ++b) pp[a][b] = *reinterpret_cast<const ulonglong2*>(&p1[(a * X1B + b) * FPD + 4 * f2]);
#pragma unroll
        for (int fi = 0; fi < 2; ++fi) {
          ulonglong2 wt[4];                // .x = channels (4 cg, 4 cg + 1), .y = (4 cg + 2, 4 cg + 3)
#pragma unroll
          for (int tap = 0; tap < 4; ++tap)
            wt[tap] = *reinterpret_cast<const ulonglong2*>(&s_w2[(tap * F + 2 * f2 + fi) * C + 4 * cg]);
#pragma unroll
          for (int dt = 0; dt < 2; ++dt)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx)
#pragma unroll
              for (int tap = 0; tap < 4; ++tap) {
                const ulonglong2 q = pp[dt + (tap >> 1)][dx + (tap & 1)];
                const f32x2 a = fi == 0 ? q.x : q.y;
                acc[dt][dx][0] = fma2(a, wt[tap].x, acc[dt][dx][0]);
                acc[dt][dx][1] = fma2(a, wt[tap].y, acc[dt][dx][1]);
              }
        }
      }
      float o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const f32x2 a = (&acc[0][0][0])[2 * i + (j >> 1)];
          v[i] = (j & 1) ? hi2(a) : lo2(a);
        }
        const float best = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
        o[j] = fmaxf(__fadd_rn(best, s_b2[4 * cg + j]), 0.f);
      }
      float* dst = p.feat + static_cast<size_t>(n0 + ch) * T2 * X2 * C + (t2 * X2 + x2) * C + 4 * cg;
      __stcs(reinterpret_cast<float4*>(dst), make_float4(o[0], o[1], o[2], o[3]));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Tensor-core variant (F in {8, 16, 32}, T and X multiples of 4): phase 1 as above on the FMA pipe; the second
// convolution -- over a batch a [positions] x [4 taps x F] x [2F] contraction, more than half of the front-end's
// multiply-adds -- as an IMPLICIT GEMM on tcgen05 with the 3xTF32 split of gemm_tc.cu (hi.lo + lo.hi + hi.hi, fp32
// accumulation in TMEM; K = 4F <= 128 per output, far below the 128-deep flush the long products need).
//
//   A operand = the pooled conv-1 plane P1 itself, never an im2col copy.  Phase 1 writes P1 (split into TF32 hi / lo)
//   into shared memory as  A[part][f / 4][row][f % 4],  row = chain * R1 + t1 * (X1 + 1) + x1  over a (T1+1) x (X1+1)
//   grid whose last row / column stay zero (TF's 'same' padding of the 2-wide kernel).  That is the UMMA K-major
//   SWIZZLE_NONE layout with rows 16 B apart (SBO = 128 B, LBO = ROWS * 16 B), so the operand of tap (a, b) is the
//   SAME buffer with the start address advanced by a * (X1 + 1) + b rows: four taps = four accumulating MMAs per
//   K step, no data movement.  Rows of the padding column / of the tail produce values nobody reads.
//   B operand = w2[tap][kd = 0] as [part][tap][f / 4][c][f % 4], split once per CTA.
//   D: M tile = 128 consecutive rows (chains back to back), N = 2F columns of TMEM per tile.
//   Epilogue: tcgen05.ld (a thread owns one row), rows staged through shared memory (XOR-swizzled float4 groups), then
//   the 2x2 max-pool + bias + ReLU of the generic kernel and the feature store.
// One elected thread issues the MMAs and commits them to an mbarrier the CTA waits on; with two CTAs per SM one CTA's
// MMAs / epilogue run under the other's phase 1.
template <int F, int PF, int TH>
__global__ void __launch_bounds__(TH, 512 / TH) conv3d_tc_kernel(const ConvParams p, const int CH, const int MT,
                                                                  const int ROWS) {
  extern __shared__ __align__(128) float sm[];
  constexpr int C = 2 * F, CG = C / 4, FH = F / 2, G = TH / FH, KC = F / 4;
  constexpr uint32_t kIdesc = umma_idesc_tf32(128, C);
  const int T = p.T, X = p.X;
  const int T1 = T / 2, X1 = X / 2, T2 = T1 / 2, X2 = X1 / 2;
  const int D = 2 * T * X, XB = X + 2, IN = (T + 2) * XB * 4;          // bordered plane: [t][x][mu][dup]
  const int X1B = X1 + 1, R1 = (T1 + 1) * X1B;                        // rows of one chain's padded P1 grid
  const int APART = KC * ROWS * 4, BPART = 4 * F * C;                 // floats per hi / lo part
  float* s_bw = sm;                        // [2][4][KC][C][4]
  float* s_b2 = s_bw + 2 * BPART;          // [C]
  float* s_a = s_b2 + C;                   // [2][KC][ROWS][4]   (16-byte rows: 128-byte aligned start)
  float* s_z = s_a + 2 * APART;            // [MT * 128][C], float4 groups XOR-swizzled by row
  float* s_in = s_z + MT * 128 * C;        // [CH][IN]
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_in + CH * IN);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_wait();
  pdl_trigger();

  {                                                                    // w2[kt][kx][kd = 0][f][c], hi / lo split
    constexpr int NW4 = F * C, PER = (NW4 + TH - 1) / TH;              // float4 groups; all loads of a thread in flight at once
    float4 wq[PER];
#pragma unroll
    for (int q = 0; q < PER; ++q) {
      const int i = 4 * (tid + q * TH), tap = i / (F * C), rest = i % (F * C);
      wq[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < 4 * F * C) wq[q] = __ldg(reinterpret_cast<const float4*>(p.w2 + static_cast<size_t>(tap) * 2 * F * C + rest));
    }
#pragma unroll
    for (int q = 0; q < PER; ++q) {
      const int i = 4 * (tid + q * TH), tap = i / (F * C), rest = i % (F * C), f = rest / C, c = rest % C;
      if (i >= 4 * F * C) continue;
      const float wv[4] = {wq[q].x, wq[q].y, wq[q].z, wq[q].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float hi = tf32_hi(wv[e]);
        const int dst = ((tap * KC + f / 4) * C + c + e) * 4 + (f & 3);
        s_bw[dst] = hi;
        s_bw[BPART + dst] = tf32_lo(wv[e], hi);
      }
    }
  }
  for (int i = tid; i < C; i += TH) s_b2[i] = p.b2[i];
  for (int i = tid; i < CH * IN; i += TH) s_in[i] = 0.f;    // borders stay zero for the whole kernel
  for (int i = tid; i < 2 * APART; i += TH) s_a[i] = 0.f;   // the padding row / column of every chain's grid stays zero
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(MT * C)) tmem_cols <<= 1;
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  uint32_t bar_phase = 0;

  const int fp = tid % FH, grp = tid / FH;                             // filters 2 fp, 2 fp + 1
  f32x2 w[3][3][2];
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < 3; ++b)
#pragma unroll
      for (int m = 0; m < 2; ++m)
        w[a][b][m] = pack2(p.w1[((a * 3 + b) * 2 + m) * F + 2 * fp], p.w1[((a * 3 + b) * 2 + m) * F + 2 * fp + 1]);
  const float bias1a = p.b1[2 * fp], bias1b = p.b1[2 * fp + 1];
  const int quads = D / 4;                 // float4 per chain

  float4 pre[PF];
  auto fetch = [&](int n0, int cnt) {      // links of chains n0 .. n0 + cnt - 1 -> registers
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i4 = tid + q * TH;
      const int ch = i4 / quads;
      pre[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ch < cnt)
        pre[q] = __ldcs(reinterpret_cast<const float4*>(p.in + static_cast<size_t>(n0 + ch) * D) + (i4 - ch * quads));
    }
  };
  auto stage = [&](int n0, int cnt) {      // registers -> bordered plane, mask applied, values duplicated
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i4 = tid + q * TH;
      const int ch = i4 / quads;
      if (ch >= cnt) continue;
      const int j = (i4 - ch * quads) * 4;
      float4 a = pre[q];
      const bool alt = p.row_set != nullptr && p.row_set[n0 + ch] != 0;
      const float* mask = alt ? p.mask_alt : p.mask;
      if (mask) {
        float4 m = *reinterpret_cast<const float4*>(mask + j);
        if (alt ? p.invert_alt : p.invert)
          m = make_float4(__fsub_rn(1.f, m.x), __fsub_rn(1.f, m.y), __fsub_rn(1.f, m.z), __fsub_rn(1.f, m.w));
        a = make_float4(__fmul_rn(m.x, a.x), __fmul_rn(m.y, a.y), __fmul_rn(m.z, a.z), __fmul_rn(m.w, a.w));
      }
      const int site = j >> 1, t = site / X, x = site - t * X;       // two sites (x, x + 1) of one row
      float* dst = s_in + ch * IN + ((t + 1) * XB + x + 1) * 4;
      *reinterpret_cast<float4*>(dst) = make_float4(a.x, a.x, a.y, a.y);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(a.z, a.z, a.w, a.w);
    }
  };

  // This CTA's chains: a contiguous block of N / grid (+1) chains, cut into passes of nearly equal size (<= CH chains).
  // A launch is only a few passes per CTA, so an even split decides the time (2048 chains on 296 CTAs: 7 chains = passes
  // of 3 + 2 + 2 instead of a grid-strided 3 + 3 + 3).
  const int nb = static_cast<int>(gridDim.x), bi = static_cast<int>(blockIdx.x);
  const int c_cnt = p.N / nb + (bi < p.N % nb ? 1 : 0), c_lo = bi * (p.N / nb) + min(bi, p.N % nb);
  const int npass = (c_cnt + CH - 1) / CH;
  auto pass_lo = [&](int j) { return c_lo + j * (c_cnt / npass) + min(j, c_cnt % npass); };
  if (npass > 0) {
    fetch(c_lo, pass_lo(1) - c_lo);
    __syncthreads();                       // the zero fills are complete
    stage(c_lo, pass_lo(1) - c_lo);
    if (npass > 1) fetch(pass_lo(1), pass_lo(2) - pass_lo(1));
  }
  for (int j = 0; j < npass; ++j) {
    const int n0 = pass_lo(j), nch = pass_lo(j + 1) - n0;
    __syncthreads();                       // the plane staged during the previous pass is visible; s_z is free

    // ---- phase 1: one pooled output (t1, x1) of filters (2 fp, 2 fp + 1) per item -> P1 as the TF32-split A operand
    for (int item = grp; item < nch * T1 * X1; item += G) {
      const int x1 = item % X1, rest = item / X1, t1 = rest % T1, ch = rest / T1;
      const float* in_c = s_in + ch * IN + (2 * t1 * XB + 2 * x1) * 4;   // bordered coords of (2 t1 - 1, 2 x1 - 1)
      f32x2 s0[2][2], s1[2][2];            // conv outputs at mu = 0 / mu = 1 of the four window positions
#pragma unroll
      for (int i = 0; i < 4; ++i) (&s0[0][0])[i] = 0ull, (&s1[0][0])[i] = 0ull;
#pragma unroll
      for (int r = 0; r < 4; ++r) {        // patch row r feeds window row dt through tap row a = r - dt
        ulonglong2 row[4];                 // .x = (mu0, mu0), .y = (mu1, mu1)
#pragma unroll
        for (int b = 0; b < 4; ++b) row[b] = *reinterpret_cast<const ulonglong2*>(&in_c[(r * XB + b) * 4]);
#pragma unroll
        for (int dt = 0; dt < 2; ++dt) {
          const int a = r - dt;
          if (a < 0 || a > 2) continue;
#pragma unroll
          for (int dx = 0; dx < 2; ++dx)
#pragma unroll
            for (int b = 0; b < 3; ++b) {
              s0[dt][dx] = fma2(row[dx + b].x, w[a][b][0], s0[dt][dx]);
              s0[dt][dx] = fma2(row[dx + b].y, w[a][b][1], s0[dt][dx]);
              s1[dt][dx] = fma2(row[dx + b].y, w[a][b][0], s1[dt][dx]);
            }
        }
      }
      float ba = -INFINITY, bb = -INFINITY;
      int aa = 0, ab = 0;                  // first maximum in (dt, dx, mu) order (training tape)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a0 = lo2((&s0[0][0])[i]), a1 = lo2((&s1[0][0])[i]);
        const float b0 = hi2((&s0[0][0])[i]), b1 = hi2((&s1[0][0])[i]);
        if (a0 > ba) ba = a0, aa = 2 * i;
        if (a1 > ba) ba = a1, aa = 2 * i + 1;
        if (b0 > bb) bb = b0, ab = 2 * i;
        if (b1 > bb) bb = b1, ab = 2 * i + 1;
      }
      ba = fmaxf(__fadd_rn(ba, bias1a), 0.f);
      bb = fmaxf(__fadd_rn(bb, bias1b), 0.f);
      if (p.t_p1 != nullptr) {
        const size_t at = (static_cast<size_t>(n0 + ch) * T1 * X1 + t1 * X1 + x1) * F + 2 * fp;
        *reinterpret_cast<float2*>(p.t_p1 + at) = make_float2(ba, bb);
        *reinterpret_cast<uchar2*>(p.t_a1 + at) = make_uchar2(ba > 0.f ? aa : 255, bb > 0.f ? ab : 255);
      }
      const float ha = tf32_hi(ba), hb = tf32_hi(bb);
      float* dst = s_a + (((2 * fp) >> 2) * ROWS + ch * R1 + t1 * X1B + x1) * 4 + ((2 * fp) & 3);
      *reinterpret_cast<float2*>(dst) = make_float2(ha, hb);
      *reinterpret_cast<float2*>(dst + APART) = make_float2(tf32_lo(ba, ha), tf32_lo(bb, hb));
    }
    fence_proxy_async_smem();              // P1 (generic-proxy stores) -> visible to the tensor core's reads
    __syncthreads();

    // ---- conv 2: per M tile and tap, F / 8 K steps x (hi.lo + lo.hi + hi.hi)
    if (tid == 0) {
      tc_fence_after();
      const uint32_t a_base = smem_u32(s_a), b_base = smem_u32(s_bw);
      const uint32_t a_lbo = static_cast<uint32_t>(ROWS) * 16u, b_lbo = static_cast<uint32_t>(C) * 16u;
      for (int mt = 0; mt < MT; ++mt) {
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(mt * C);
        uint32_t acc = 0;
#pragma unroll
        for (int tap = 0; tap < 4; ++tap) {
          const int shift = (tap >> 1) * X1B + (tap & 1);
#pragma unroll
          for (int ks = 0; ks < F / 8; ++ks) {
            const uint32_t a_off = (static_cast<uint32_t>(2 * ks * ROWS + mt * 128 + shift)) * 16u;
            const uint32_t b_off = (static_cast<uint32_t>((tap * KC + 2 * ks) * C)) * 16u;
            const uint64_t a_hi = umma_desc_kmajor(a_base + a_off, a_lbo, 128u);
            const uint64_t a_lo = umma_desc_kmajor(a_base + static_cast<uint32_t>(APART) * 4u + a_off, a_lbo, 128u);
            const uint64_t b_hi = umma_desc_kmajor(b_base + b_off, b_lbo, 128u);
            const uint64_t b_lo = umma_desc_kmajor(b_base + static_cast<uint32_t>(BPART) * 4u + b_off, b_lbo, 128u);
            umma_tf32(d_tmem, a_hi, b_lo, kIdesc, acc);
            umma_tf32(d_tmem, a_lo, b_hi, kIdesc, 1);
            umma_tf32(d_tmem, a_hi, b_hi, kIdesc, 1);
            acc = 1;
          }
        }
      }
      umma_commit(bar);
    }
    // under the tensor core's latency: the next pass's links into the plane (phase 1 is done with it), the pass after
    // that into registers
    if (j + 1 < npass) {
      stage(pass_lo(j + 1), pass_lo(j + 2) - pass_lo(j + 1));
      if (j + 2 < npass) fetch(pass_lo(j + 2), pass_lo(j + 3) - pass_lo(j + 2));
    }
    if (tid == 0) mbar_wait_sleep(bar, bar_phase);   // one waiter: 255 polling threads would take the other CTA's issue slots
    bar_phase ^= 1;
    __syncthreads();
    tc_fence_after();

    // ---- epilogue: TMEM -> shared (thread = row), then pool (2,2,[1]) + bias + relu -> flattened features
    for (int mt = warp >> 2; mt < MT; mt += TH / 128) {
      const int row = mt * 128 + (warp & 3) * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + static_cast<uint32_t>(mt * C);
      float* zrow = s_z + row * C;
#pragma unroll
      for (int c0 = 0; c0 < C; c0 += 32) {
        float v[32];
        if (C >= 32) tmem_ld32(taddr + c0, v);
        else tmem_ld16(taddr + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < (C >= 32 ? 8 : 4); ++g) {
          const int grp4 = (c0 / 4 + g) ^ (row & (CG - 1));
          *reinterpret_cast<float4*>(zrow + grp4 * 4) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        }
      }
    }
    tc_fence_before();
    __syncthreads();
    const int cg = tid % CG;
    for (int item = tid / CG; item < nch * T2 * X2; item += TH / CG) {
      const int x2 = item % X2, rest = item / X2, t2 = rest % T2, ch = rest / T2;
      const int r00 = ch * R1 + 2 * t2 * X1B + 2 * x2;
      float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      int arg[4] = {0, 0, 0, 0};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int row = r00 + (k >> 1) * X1B + (k & 1);
        const float4 z4 = *reinterpret_cast<const float4*>(s_z + row * C + ((cg ^ (row & (CG - 1))) * 4));
        const float z[4] = {z4.x, z4.y, z4.z, z4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (z[e] > best[e]) best[e] = z[e], arg[e] = k;
      }
      const float4 b = *reinterpret_cast<const float4*>(s_b2 + 4 * cg);
      const float o[4] = {fmaxf(__fadd_rn(best[0], b.x), 0.f), fmaxf(__fadd_rn(best[1], b.y), 0.f),
                          fmaxf(__fadd_rn(best[2], b.z), 0.f), fmaxf(__fadd_rn(best[3], b.w), 0.f)};
      const size_t at = static_cast<size_t>(n0 + ch) * T2 * X2 * C + (t2 * X2 + x2) * C + 4 * cg;
      __stcs(reinterpret_cast<float4*>(p.feat + at), make_float4(o[0], o[1], o[2], o[3]));
      if (p.t_a2 != nullptr)
        *reinterpret_cast<uchar4*>(p.t_a2 + at) = make_uchar4(o[0] > 0.f ? arg[0] : 255, o[1] > 0.f ? arg[1] : 255,
                                                              o[2] > 0.f ? arg[2] : 255, o[3] > 0.f ? arg[3] : 255);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, tmem_cols);
}

// -1: the shape does not fit (shared memory); the FFMA kernels below take it
template <int F>
int launch_tc(const ConvParams& p, int sms, int max_smem, cudaStream_t stream) {
  constexpr int TH = 256, C = 2 * F;
  const auto misaligned = [](const void* a) { return (reinterpret_cast<uintptr_t>(a) & 15u) != 0; };
  if (p.T % 4 != 0 || p.X % 4 != 0 || misaligned(p.in) || misaligned(p.feat) || misaligned(p.mask) ||
      misaligned(p.mask_alt) || misaligned(p.w2))
    return -1;
  const int T1 = p.T / 2, X1 = p.X / 2, R1 = (T1 + 1) * (X1 + 1), D = 2 * p.T * p.X;
  // chains per pass: at most two M tiles of rows and 4 float4 of links per thread
  const int CH = std::max(1, std::min(256 / R1, 4 * TH * 4 / D));
  const int MT = (CH * R1 + 127) / 128;
  int ROWS = MT * 128 + X1 + 2;                                        // the last tile's taps read up to X1 + 2 rows further
  while ((ROWS * 4) % 32 != 8) ++ROWS;                                 // K chunks land 8 banks apart
  const size_t floats = 2u * 4 * F * C + C + 2u * (F / 4) * ROWS * 4 + static_cast<size_t>(MT) * 128 * C +
                        static_cast<size_t>(CH) * (p.T + 2) * (p.X + 2) * 4;
  const size_t smem = sizeof(float) * floats + 16;
  if (smem > static_cast<size_t>(max_smem) || MT * C > 512) return -1;
  const int pf = (CH * D / 4 + TH - 1) / TH;
  const int per_sm = std::max(1, std::min(static_cast<int>((200 * 1024) / (smem + 1024)), 512 / TH));
  const int grid = std::max(1, std::min(p.N, sms * per_sm));           // every CTA gets N / grid (+1) chains
  auto go = [&](auto kern) -> int {
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
      if (e != cudaSuccess) {
        set_error("cudaFuncSetAttribute(conv3d_tc_kernel): %s", cudaGetErrorString(e));
        return L2HMC_ERR_CUDA;
      }
    }
    launch_pdl(kern, dim3(grid), dim3(TH), smem, stream, p, CH, MT, ROWS);
    return check_launch("conv3d_tc_kernel");
  };
  if (pf <= 1) return go(conv3d_tc_kernel<F, 1, TH>);
  if (pf <= 2) return go(conv3d_tc_kernel<F, 2, TH>);
  if (pf <= 4) return go(conv3d_tc_kernel<F, 4, TH>);
  return -1;
}

template <int F, int PF, int CH, int TH>
int launch_tiled_pf(const ConvParams& p, int sms, size_t smem, cudaStream_t stream) {
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(conv3d_tiled_kernel<F, PF, CH, TH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv3d_tiled_kernel): %s", cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  const int per_sm = max(1, min(static_cast<int>((200 * 1024) / (smem + 1024)), 512 / TH));
  const int grid = max(1, min((p.N + CH - 1) / CH, sms * per_sm));
  conv3d_tiled_kernel<F, PF, CH, TH><<<grid, TH, smem, stream>>>(p);
  return check_launch("conv3d_tiled_kernel");
}

template <int F, int CH, int TH>
int launch_tiled_ch(const ConvParams& p, int sms, size_t smem, cudaStream_t stream) {
  const int pf = (CH * 2 * p.T * p.X / 4 + TH - 1) / TH;
  if (pf <= 1) return launch_tiled_pf<F, 1, CH, TH>(p, sms, smem, stream);
  if (pf <= 2) return launch_tiled_pf<F, 2, CH, TH>(p, sms, smem, stream);
  if (pf <= 4) return launch_tiled_pf<F, 4, CH, TH>(p, sms, smem, stream);
  return -1;
}

// -1: shape not covered (odd extents, misaligned arrays, too large for shared memory) -> generic kernel
template <int F>
int launch_tiled(const ConvParams& p, int sms, int max_smem, cudaStream_t stream) {
  const auto misaligned = [](const void* a) { return (reinterpret_cast<uintptr_t>(a) & 15u) != 0; };
  if (p.T % 4 != 0 || p.X % 4 != 0 || misaligned(p.in) || misaligned(p.feat) || misaligned(p.mask) ||
      misaligned(p.mask_alt))
    return -1;
  const int T1 = p.T / 2, X1 = p.X / 2;
  const auto smem_for = [&](int ch) {
    return sizeof(float) * (4 * F * 2 * F + 2 * F + static_cast<size_t>(ch) * (p.T + 2) * (p.X + 2) * 4 +
                            static_cast<size_t>(ch) * (T1 + 1) * (X1 + 1) * (2 * F + 8));
  };
  // two chains per pass while two CTAs still fit an SM, one chain per pass for the large lattices
  // (four 128-thread CTAs per SM with one chain each measured the same: the FMA pipe is the limit)
  if (smem_for(2) <= 100 * 1024) return launch_tiled_ch<F, 2, 256>(p, sms, smem_for(2), stream);
  if (smem_for(1) <= static_cast<size_t>(max_smem)) return launch_tiled_ch<F, 1, 256>(p, sms, smem_for(1), stream);
  return -1;
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_conv3d_num_features(int T, int X, int F) {
  const int T1 = (T + 1) / 2, X1 = (X + 1) / 2;
  return ((T1 + 1) / 2) * ((X1 + 1) / 2) * 2 * F;
}

extern "C" int l2hmc_conv3d_features(const l2hmc_conv3d_weights* w, const float* in, const float* mask,
                                     int mask_invert, int N, int T, int X, float* feat, void* stream) {
  return l2hmc_conv3d_features_sets(w, in, mask, mask_invert, nullptr, 0, nullptr, N, T, X, feat, stream);
}

extern "C" int l2hmc_conv3d_features_sets(const l2hmc_conv3d_weights* w, const float* in, const float* mask,
                                          int mask_invert, const float* mask_alt, int mask_invert_alt,
                                          const uint8_t* row_set, int N, int T, int X, float* feat, void* stream) {
  return l2hmc_conv3d_features_tape(w, in, mask, mask_invert, mask_alt, mask_invert_alt, row_set, N, T, X, feat, nullptr,
                                    stream);
}

extern "C" int l2hmc_conv3d_tape_elems(int T, int X, int F) { return ((T + 1) / 2) * ((X + 1) / 2) * F; }

extern "C" int l2hmc_conv3d_features_tape(const l2hmc_conv3d_weights* w, const float* in, const float* mask,
                                          int mask_invert, const float* mask_alt, int mask_invert_alt,
                                          const uint8_t* row_set, int N, int T, int X, float* feat,
                                          const l2hmc_conv3d_tape* tape, void* stream) {
  L2HMC_REQUIRE(w, "l2hmc_conv3d_features: null weights");
  L2HMC_REQUIRE(!tape || (tape->p1 && tape->arg1 && tape->arg2), "l2hmc_conv3d_features_tape: null tape array");
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(w->w1 && w->b1 && w->w2 && w->b2 && in && feat, "l2hmc_conv3d_features: null pointer");
  L2HMC_REQUIRE(N >= 0 && T > 0 && X > 0 && w->F > 0, "l2hmc_conv3d_features: bad sizes N=%d T=%d X=%d F=%d", N, T, X,
                w->F);
  const int F = w->F, C = 2 * F, D = 2 * T * X, T1 = (T + 1) / 2, X1 = (X + 1) / 2;
  const size_t smem = sizeof(float) * (((18 * F + F + 4 * F * C + C + 3) & ~3) + ((D + 3) & ~3) + T1 * X1 * F);
  int dev = 0, max_smem = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  ConvParams p{w->w1, w->b1, w->w2, w->b2, in, mask, mask_invert, mask_alt, mask_invert_alt, row_set, N, T, X, F, feat,
               tape ? tape->p1 : nullptr, tape ? tape->arg1 : nullptr, tape ? tape->arg2 : nullptr};
  int rc = -1;
  // conv 2 on the tensor cores unless the shape does not fit or L2HMC_CONV_PATH=ffma asks for the FMA-pipe kernels
  static const bool ffma_only = [] { const char* e = std::getenv("L2HMC_CONV_PATH"); return e && std::string(e) == "ffma"; }();
  if (!ffma_only) {
    if (F == 8) rc = launch_tc<8>(p, sms, max_smem, static_cast<cudaStream_t>(stream));
    else if (F == 16) rc = launch_tc<16>(p, sms, max_smem, static_cast<cudaStream_t>(stream));
    else if (F == 32) rc = launch_tc<32>(p, sms, max_smem, static_cast<cudaStream_t>(stream));
    if (rc >= 0) return rc;
  }
  if (tape == nullptr) {                  // (the all-FFMA2 tiled kernel does not record the training tape)
    if (F == 8) rc = launch_tiled<8>(p, sms, max_smem, static_cast<cudaStream_t>(stream));
    else if (F == 16) rc = launch_tiled<16>(p, sms, max_smem, static_cast<cudaStream_t>(stream));
    else if (F == 32) rc = launch_tiled<32>(p, sms, max_smem, static_cast<cudaStream_t>(stream));
  }
  if (rc >= 0) return rc;                 // -1: not covered by the tiled kernel, use the generic one
  if (smem > static_cast<size_t>(max_smem)) {
    set_error("l2hmc_conv3d_features: %dx%d lattice with F=%d needs %zu B of shared memory (limit %d)", T, X, F, smem,
              max_smem);
    return L2HMC_ERR_UNSUPPORTED;
  }
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(conv3d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv3d_kernel): %s", cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  int per_sm = static_cast<int>((200 * 1024) / (smem + 1024));
  per_sm = max(1, min(per_sm, 8));
  const int grid = max(1, min(N, sms * per_sm));
  conv3d_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("conv3d_kernel");
}
