// Backward pass of the ConvNet3D front-end (one input stream) FROM THE FORWARD PASS'S ROUTING TAPE:
//   (N,T,X,2,1) -> Conv3D(F,(3,3,2),same,relu) -> MaxPool3D(2,2,same) -> Conv3D(2F,(2,2,2),same,relu)
//               -> MaxPool3D(2,2,same) -> flatten                       (reference l2hmc/network/conv_net.py:247-260)
// The reference gets these gradients from tf.gradients (gauge_model.py:823), which keeps every activation.  The adjoint
// needs far less: the pooled conv-1 plane P1 (the conv-2 weight gradient's operand) and, per pooling window, WHICH
// candidate was the maximum -- l2hmc_conv3d_tape, 5.5 KB per chain and call at 16x16, written by
// l2hmc_conv3d_features_tape.  With it
//   * nothing is recomputed (conv3d_bwd_kernel recomputes both convolutions: a third of its arithmetic), and
//   * every product runs over the NON-ZEROS only: a max-pool passes its gradient to one candidate per window, so the
//     gradient of the conv-2 pre-activation has one non-zero per 2x2 window and channel (1/4 dense) and that of the
//     conv-1 pre-activation one per 2x2x2 window and filter (1/8 dense):
//       conv-2 weight gradient   thread = (tap, channel, 8 inputs), loops over the T2*X2 windows     (1/4 of the dense FMAs)
//       conv-2 input gradient    thread = (window, input filter): the window's non-zero lands on a 3x3 patch held in
//                                registers; patches meet in shared memory and are added in a fixed order   (1/4)
//       conv-1 weight gradient   thread = (filter, slab of windows): per non-zero the 3x3x2 input patch around it (from a
//                                zero-bordered plane: no bounds logic) into 18 register accumulators           (1/8)
//       conv-1 input gradient    the non-zeros are scattered into a dense conv-1 gradient plane (each has its own slot) and
//                                gathered per site with 128-bit loads over the filters -- measured cheaper than any
//                                per-non-zero gather: decoding a location costs more than the seven FMAs on zeros it saves
//   A ReLU-dead maximum (pooled value 0) is code 255 on the tape and carries no gradient.
// One chain per CTA pass; weight gradients accumulate into the CTA's row of `partial` exactly as in conv3d_bwd.cu
// (deterministic, no atomics).  Same gradients as the recompute kernel up to summation order.
#include "conv3d_bwd.cuh"

namespace l2hmc {
namespace {

constexpr int kThreads = 256;

struct TapeBwdParams {
  const float *w1, *w2;
  int F;
  const float* in;
  const float* mask;
  int invert;
  const float* mask_alt;
  int invert_alt;
  const uint8_t* row_set;
  int N, T, X;
  const float* g_feat;
  float* g_in;
  float* partial;
  int psize;
  const float* t_p1;
  const uint8_t* t_a1;
  const uint8_t* t_a2;
  int S;        // slabs of pooled positions in the conv-1 weight gradient
  int dense_in; // the dense conv-1 gradient plane fits shared memory (else the input gradient gathers per non-zero)
};

// FT > 0: filter count known at compile time (multiple of 8: vector loads, unrolled channel loops); FT = 0: any F.
// XT > 0: the lattice extent X too (the reference sizes its filters F = space_size, gauge_dynamics.py:121-143, so X == F
// is the usual case): the (t, x) decodes become shifts.
template <int FT, int XT>
__global__ void __launch_bounds__(kThreads, FT == 32 ? 2 : 4) conv3d_bwd_tape_kernel(const TapeBwdParams p) {
  extern __shared__ __align__(16) float sm[];
  constexpr bool kFast = FT > 0;
  const int T = p.T, X = XT > 0 ? XT : p.X, F = kFast ? FT : p.F, C = 2 * F;
  const int T1 = (T + 1) / 2, X1 = (X + 1) / 2, T2 = (T1 + 1) / 2, X2 = (X1 + 1) / 2;
  const int D = 2 * T * X, P = T1 * X1, W = T2 * X2, K = W * C;
  const int PS = kFast ? F + 4 : F + 1;   // p1 row stride (lanes that read different rows hit different banks)
  const int XB = X + 2, INB = (T + 2) * XB * 2;              // zero-bordered input plane [T+2][X+2][2]
  const int FP = kFast ? F + 4 : F + 1;                      // row stride of the dense conv-1 gradient plane
  float* inb = sm;                        // bordered masked input (the border stays zero for the whole kernel)
  float* mk_s = inb + ((INB + 3) & ~3);   // [T][X][2] the mask actually applied (1 if none)
  float* w1s = mk_s + ((D + 3) & ~3);     // [3][3][2][F]
  float* p1s = w1s + ((18 * F + 3) & ~3); // [P][PS] pooled conv-1 output; later gp1 [P][F]: its (routed) gradient
  // union: {gq, w2t, stage} until the conv-2 products are done, then the dense conv-1 gradient plane gc1 [T][X][2][FP]
  float* un = p1s + ((P * PS + 3) & ~3);
  float* gq = un;                         // [W][C] gradient of the conv-2 pre-activation at the window's maximum
  float* w2t = gq + W * C;                // [4 taps][C][F]  (kd = 0), input filter fastest
  float* stage = w2t + 4 * C * F;         // [W][9][F] input-gradient patches; later the conv-1 slab sums [S][18][F]
  float* gc1 = un;
  const int un_floats = max(W * C + 4 * C * F + max(W * 9 * F, p.S * 18 * F), p.dense_in ? D * FP : 0);
  uint8_t* aq = reinterpret_cast<uint8_t*>(un + ((un_floats + 3) & ~3));   // [W][C] dt*2 + dx of the maximum
  // [P][F] where the first pool's maximum sits in the conv-1 output: t << 8 | x << 1 | mu (0xFFFF = ReLU-dead), decoded
  // once per chain so that the loops below only shift and mask
  uint16_t* loc = reinterpret_cast<uint16_t*>(aq + ((W * C + 15) & ~15));
  float* gp1 = p1s;
  const int tid = threadIdx.x;
  for (int i = tid; i < 18 * F; i += kThreads) w1s[i] = p.w1[i];
  for (int i = tid; i < INB; i += kThreads) inb[i] = 0.f;
  float* part = p.partial + static_cast<size_t>(blockIdx.x) * p.psize;
  const int off_b1 = 18 * F, off_w2 = 19 * F, off_b2 = 19 * F + 8 * F * C;
  // Fast path: this thread's weight-gradient sums stay in REGISTERS over all the chains of the CTA and reach the partial
  // row once at the end (a global read-modify-write per chain was the kernel's long-scoreboard stall).
  constexpr int NI2 = kFast ? (4 * 2 * FT * (FT / 8) + kThreads - 1) / kThreads : 1;   // conv-2 items per thread
  constexpr int NI1 = kFast ? (18 * FT + kThreads - 1) / kThreads : 1;                 // conv-1 kernel entries per thread
  float acc_w2[NI2][8], acc_w1[NI1], acc_b2 = 0.f, acc_b1 = 0.f;
#pragma unroll
  for (int q = 0; q < NI2; ++q)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_w2[q][j] = 0.f;
#pragma unroll
  for (int q = 0; q < NI1; ++q) acc_w1[q] = 0.f;

  for (int n = blockIdx.x; n < p.N; n += gridDim.x) {
    __syncthreads();                       // previous chain done with every buffer (weights visible on the first pass)
    // ---- load: masked input, P1 and the routing codes; scatter the feature gradient to the windows' maxima
    const bool alt = p.row_set != nullptr && p.row_set[n] != 0;
    const float* mask = alt ? p.mask_alt : p.mask;
    const int invert = alt ? p.invert_alt : p.invert;
    const float* src = p.in + static_cast<size_t>(n) * D;
    for (int i = tid; i < D; i += kThreads) {
      const float m = mask == nullptr ? 1.f : (invert ? 1.f - mask[i] : mask[i]);
      const int site = i >> 1, t = site / X, x = site - t * X;
      mk_s[i] = m;
      inb[((t + 1) * XB + x + 1) * 2 + (i & 1)] = m * src[i];
    }
    for (int i = tid; i < 4 * F * C; i += kThreads) {               // (L2-resident; the union was the previous chain's gc1)
      const int co = i % C, ci = (i / C) % F, kk = i / (C * F);       // kk = kt*2 + kx; source [kt][kx][kd = 0][ci][co]
      w2t[(kk * C + co) * F + ci] = p.w2[((kk * 2 + 0) * F + ci) * C + co];
    }
    const float* tp1 = p.t_p1 + static_cast<size_t>(n) * P * F;
    const uint8_t* ta1 = p.t_a1 + static_cast<size_t>(n) * P * F;
    for (int i = tid; i < P * F; i += kThreads) {
      const int r = i / F, t1 = r / X1, x1 = r - t1 * X1, a = ta1[i];
      p1s[r * PS + i % F] = tp1[i];
      loc[i] = a == 255 ? 0xFFFF
                        : static_cast<uint16_t>(((2 * t1 + (a >> 2)) << 8) | ((2 * x1 + ((a >> 1) & 1)) << 1) | (a & 1));
    }
    const float* gf = p.g_feat + static_cast<size_t>(n) * K;
    const uint8_t* ta2 = p.t_a2 + static_cast<size_t>(n) * K;
    for (int i = tid; i < K; i += kThreads) {
      const int a = ta2[i];
      gq[i] = a == 255 ? 0.f : gf[i];
      aq[i] = a == 255 ? 0 : static_cast<uint8_t>(a);
    }
    __syncthreads();

    // ---- conv-2: bias and weight gradients over the windows' non-zeros
    for (int co = tid; co < C; co += kThreads) {
      float s = 0.f;
      for (int w = 0; w < W; ++w) s += gq[w * C + co];
      if (kFast) acc_b2 += s;        // C <= 64 < kThreads: one channel per thread
      else part[off_b2 + co] += s;
    }
    if (kFast) {       // thread = (tap, output channel, 8 input filters)
#pragma unroll
      for (int q = 0; q < NI2; ++q) {
        const int i = tid + q * kThreads;
        if (i >= 4 * C * (F / 8)) break;
        const int co = i % C, ci0 = ((i / C) % (F / 8)) * 8, kk = i / (C * (F / 8)), kt = kk >> 1, kx = kk & 1;
        float s8[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) s8[j] = 0.f;
        for (int t2 = 0; t2 < T2; ++t2)
          for (int x2 = 0; x2 < X2; ++x2) {
            const int w = t2 * X2 + x2, a = aq[w * C + co];
            const float g = gq[w * C + co];
            const int t = 2 * t2 + (a >> 1) + kt, x = 2 * x2 + (a & 1) + kx;
            if (t >= T1 || x >= X1) continue;           // the tap reads conv-2's end padding
            const float* ap = p1s + (t * X1 + x) * PS + ci0;
            const float4 a0 = *reinterpret_cast<const float4*>(ap), a1 = *reinterpret_cast<const float4*>(ap + 4);
            s8[0] = fmaf(a0.x, g, s8[0]), s8[1] = fmaf(a0.y, g, s8[1]), s8[2] = fmaf(a0.z, g, s8[2]);
            s8[3] = fmaf(a0.w, g, s8[3]), s8[4] = fmaf(a1.x, g, s8[4]), s8[5] = fmaf(a1.y, g, s8[5]);
            s8[6] = fmaf(a1.z, g, s8[6]), s8[7] = fmaf(a1.w, g, s8[7]);
          }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc_w2[q][j] += s8[j];
      }
    } else {
      for (int i = tid; i < 4 * F * C; i += kThreads) {
        const int co = i % C, ci = (i / C) % F, kk = i / (C * F), kt = kk >> 1, kx = kk & 1;
        float s = 0.f;
        for (int t2 = 0; t2 < T2; ++t2)
          for (int x2 = 0; x2 < X2; ++x2) {
            const int w = t2 * X2 + x2, a = aq[w * C + co];
            const int t = 2 * t2 + (a >> 1) + kt, x = 2 * x2 + (a & 1) + kx;
            if (t >= T1 || x >= X1) continue;
            s = fmaf(p1s[(t * X1 + x) * PS + ci], gq[w * C + co], s);
          }
        part[off_w2 + ((kk * 2 + 0) * F + ci) * C + co] += s;
      }
    }
    // ---- conv-2 input gradient: a window's non-zero (at offset (dt, dx)) reaches the 3x3 patch (dt + kt, dx + kx)
    for (int i = tid; i < W * F; i += kThreads) {
      const int ci = i % F, w = i / F;
      float a00 = 0.f, a01 = 0.f, a02 = 0.f, a10 = 0.f, a11 = 0.f, a12 = 0.f, a20 = 0.f, a21 = 0.f, a22 = 0.f;
      const float* wt = w2t + ci;
#pragma unroll 4
      for (int co = 0; co < C; ++co) {
        const float g = gq[w * C + co];
        const int a = aq[w * C + co];
        const float u00 = wt[(0 * C + co) * F], u01 = wt[(1 * C + co) * F], u10 = wt[(2 * C + co) * F],
                    u11 = wt[(3 * C + co) * F];
        switch (a) {       // uniform over the lanes of a window (they differ in ci only)
          case 0: a00 = fmaf(u00, g, a00), a01 = fmaf(u01, g, a01), a10 = fmaf(u10, g, a10), a11 = fmaf(u11, g, a11); break;
          case 1: a01 = fmaf(u00, g, a01), a02 = fmaf(u01, g, a02), a11 = fmaf(u10, g, a11), a12 = fmaf(u11, g, a12); break;
          case 2: a10 = fmaf(u00, g, a10), a11 = fmaf(u01, g, a11), a20 = fmaf(u10, g, a20), a21 = fmaf(u11, g, a21); break;
          default: a11 = fmaf(u00, g, a11), a12 = fmaf(u01, g, a12), a21 = fmaf(u10, g, a21), a22 = fmaf(u11, g, a22); break;
        }
      }
      float* st = stage + (w * 9) * F + ci;
      st[0] = a00, st[F] = a01, st[2 * F] = a02, st[3 * F] = a10, st[4 * F] = a11, st[5 * F] = a12;
      st[6 * F] = a20, st[7 * F] = a21, st[8 * F] = a22;
    }
    __syncthreads();
    // ---- the patches of the (up to four) windows that reach a position, in fixed order; pool-1 routing: the gradient
    //      belongs to the window's maximum (loc), none when it is ReLU-dead
    for (int i = tid; i < P * F; i += kThreads) {
      const int ci = i % F, r = i / F, t = r / X1, x = r - t * X1;
      float s = 0.f;
#pragma unroll
      for (int jt = 0; jt < 2; ++jt) {
        const int t2 = (t >> 1) - jt, ot = t - 2 * t2;
        if (t2 < 0 || ot > 2) continue;
#pragma unroll
        for (int jx = 0; jx < 2; ++jx) {
          const int x2 = (x >> 1) - jx, ox = x - 2 * x2;
          if (x2 < 0 || ox > 2) continue;
          s += stage[((t2 * X2 + x2) * 9 + ot * 3 + ox) * F + ci];
        }
      }
      gp1[i] = loc[i] == 0xFFFF ? 0.f : s;
    }
    __syncthreads();
    // ---- conv-1: bias and weight gradients over the windows' non-zeros (slab sums meet in `stage`, added in slab order)
    for (int f = tid; f < F; f += kThreads) {
      float s = 0.f;
      for (int r = 0; r < P; ++r) s += gp1[r * F + f];
      if (kFast) acc_b1 += s;
      else part[off_b1 + f] += s;
    }
    // thread = (filter, slab of windows): the 3x3x2 input patch around each non-zero, 18 accumulators in registers
    const int S = p.S, per = (P + S - 1) / S;
    for (int i = tid; i < S * F; i += kThreads) {
      const int f = i % F, slab = i / F;
      float acc[9][2];
#pragma unroll
      for (int j = 0; j < 9; ++j) acc[j][0] = acc[j][1] = 0.f;
      const int r_hi = min(P, (slab + 1) * per);
      for (int r = slab * per; r < r_hi; ++r) {
        const unsigned l = loc[r * F + f];
        if (l == 0xFFFFu) continue;
        const float g = gp1[r * F + f];
        const bool mu1 = (l & 1u) != 0;
        // bordered coordinates of input (t + kt - 1, x + kx - 1) are (t + kt, x + kx)
        const float* base = inb + ((l >> 8) * XB + ((l >> 1) & 127u)) * 2;
#pragma unroll
        for (int kt = 0; kt < 3; ++kt)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const float2 iv = *reinterpret_cast<const float2*>(base + (kt * XB + kx) * 2);
            // output mu = 0: kd = 0 reads input mu = 0, kd = 1 reads input mu = 1; output mu = 1: kd = 0 reads input mu = 1
            acc[kt * 3 + kx][0] = fmaf(mu1 ? iv.y : iv.x, g, acc[kt * 3 + kx][0]);
            acc[kt * 3 + kx][1] = fmaf(mu1 ? 0.f : iv.y, g, acc[kt * 3 + kx][1]);
          }
      }
#pragma unroll
      for (int j = 0; j < 9; ++j) {
        stage[(slab * 18 + 2 * j) * F + f] = acc[j][0];
        stage[(slab * 18 + 2 * j + 1) * F + f] = acc[j][1];
      }
    }
    __syncthreads();
    if (kFast) {
#pragma unroll
      for (int k = 0; k < NI1; ++k) {
        const int i = tid + k * kThreads;
        if (i >= 18 * F) break;
        float s = 0.f;
        for (int q = 0; q < S; ++q) s += stage[q * 18 * F + i];
        acc_w1[k] += s;
      }
    } else {
      for (int i = tid; i < 18 * F; i += kThreads) {
        float s = 0.f;
        for (int q = 0; q < S; ++q) s += stage[q * 18 * F + i];
        part[i] += s;
      }
    }
    __syncthreads();
    // ---- conv-1 input gradient: scatter the non-zeros into the dense plane gc1 (over the dead union), gather per site
    float* dst = p.g_in + static_cast<size_t>(n) * D;
    if (!p.dense_in) {       // the plane does not fit (32x32 with F = 32): every site gathers from the <= 4 windows around it
      for (int site = tid; site < T * X; site += kThreads) {
        const int tt = site / X, xx = site - tt * X;
        float s0 = 0.f, s1 = 0.f;
        const int t1lo = max(tt - 1, 0) >> 1, t1hi = min(tt + 1, T - 1) >> 1;
        const int x1lo = max(xx - 1, 0) >> 1, x1hi = min(xx + 1, X - 1) >> 1;
        for (int t1 = t1lo; t1 <= t1hi; ++t1)
          for (int x1 = x1lo; x1 <= x1hi; ++x1) {
            const int r = t1 * X1 + x1;
#pragma unroll 4
            for (int f = 0; f < F; ++f) {
              const unsigned l = loc[r * F + f];
              const int kt = tt + 1 - static_cast<int>(l >> 8), kx = xx + 1 - static_cast<int>((l >> 1) & 127u);
              if (l == 0xFFFFu || static_cast<unsigned>(kt) > 2u || static_cast<unsigned>(kx) > 2u) continue;
              const float g = gp1[r * F + f];
              const float* wk = w1s + ((kt * 3 + kx) * 2) * F + f;
              if ((l & 1u) == 0) {      // output mu = 0 feeds input mu = 0 (kd = 0) and input mu = 1 (kd = 1)
                s0 = fmaf(wk[0], g, s0);
                s1 = fmaf(wk[F], g, s1);
              } else {                  // output mu = 1 feeds input mu = 1 (kd = 0)
                s1 = fmaf(wk[0], g, s1);
              }
            }
          }
        dst[2 * site] += mk_s[2 * site] * s0;
        dst[2 * site + 1] += mk_s[2 * site + 1] * s1;
      }
      continue;
    }
    for (int i = tid; i < D * FP; i += kThreads) gc1[i] = 0.f;
    __syncthreads();
    for (int i = tid; i < P * F; i += kThreads) {
      const unsigned l = loc[i];
      if (l != 0xFFFFu) gc1[((((l >> 8) * X + ((l >> 1) & 127u)) * 2 + (l & 1u))) * FP + i % F] = gp1[i];
    }
    __syncthreads();
    for (int site = tid; site < T * X; site += kThreads) {
      // mu = 0 gets tap kd = 0 of output mu = 0; mu = 1 gets kd = 0 of output mu = 1 and kd = 1 of output mu = 0
      const int tt = site / X, xx = site - tt * X;
      float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
      for (int kt = 0; kt < 3; ++kt) {
        const int t = tt - kt + 1;
        if (t < 0 || t >= T) continue;
        for (int kx = 0; kx < 3; ++kx) {
          const int x = xx - kx + 1;
          if (x < 0 || x >= X) continue;
          const float* g = gc1 + ((t * X + x) * 2) * FP;
          const float* w = w1s + ((kt * 3 + kx) * 2) * F;
          if (kFast) {
#pragma unroll
            for (int f4 = 0; f4 < F; f4 += 4) {
              const float4 w0 = *reinterpret_cast<const float4*>(w + f4);
              const float4 w1 = *reinterpret_cast<const float4*>(w + F + f4);
              const float4 g0 = *reinterpret_cast<const float4*>(g + f4);
              const float4 g1 = *reinterpret_cast<const float4*>(g + FP + f4);
              s0[0] = fmaf(w0.x, g0.x, s0[0]), s0[1] = fmaf(w0.y, g0.y, s0[1]);
              s0[2] = fmaf(w0.z, g0.z, s0[2]), s0[3] = fmaf(w0.w, g0.w, s0[3]);
              s1[0] = fmaf(w0.x, g1.x, s1[0]), s1[1] = fmaf(w0.y, g1.y, s1[1]);
              s1[2] = fmaf(w0.z, g1.z, s1[2]), s1[3] = fmaf(w0.w, g1.w, s1[3]);
              s1[0] = fmaf(w1.x, g0.x, s1[0]), s1[1] = fmaf(w1.y, g0.y, s1[1]);
              s1[2] = fmaf(w1.z, g0.z, s1[2]), s1[3] = fmaf(w1.w, g0.w, s1[3]);
            }
          } else {
            for (int f = 0; f < F; ++f) {
              s0[0] = fmaf(w[f], g[f], s0[0]);
              s1[0] = fmaf(w[f], g[FP + f], s1[0]);
              s1[1] = fmaf(w[F + f], g[f], s1[1]);
            }
          }
        }
      }
      dst[2 * site] += mk_s[2 * site] * ((s0[0] + s0[1]) + (s0[2] + s0[3]));
      dst[2 * site + 1] += mk_s[2 * site + 1] * ((s1[0] + s1[1]) + (s1[2] + s1[3]));
    }
  }
  if (kFast) {       // the CTA's weight-gradient sums -> its row of `partial`
    if (tid < C) part[off_b2 + tid] += acc_b2;
    if (tid < F) part[off_b1 + tid] += acc_b1;
#pragma unroll
    for (int q = 0; q < NI2; ++q) {
      const int i = tid + q * kThreads;
      if (i >= 4 * C * (F / 8)) break;
      const int co = i % C, ci0 = ((i / C) % (F / 8)) * 8, kk = i / (C * (F / 8));
#pragma unroll
      for (int j = 0; j < 8; ++j) part[off_w2 + ((kk * 2 + 0) * F + ci0 + j) * C + co] += acc_w2[q][j];
    }
#pragma unroll
    for (int k = 0; k < NI1; ++k) {
      const int i = tid + k * kThreads;
      if (i < 18 * F) part[i] += acc_w1[k];
    }
  }
}

}  // namespace

int conv3d_backward_tape(const l2hmc_conv3d_bwd_args* a, int rows, int psize, cudaStream_t stream) {
  const int F = a->w->F, T = a->T, X = a->X, C = 2 * F;
  const int T1 = (T + 1) / 2, X1 = (X + 1) / 2, T2 = (T1 + 1) / 2, X2 = (X1 + 1) / 2;
  const int D = 2 * T * X, P = T1 * X1, W = T2 * X2;
  const bool fast = F == 8 || F == 16 || F == 32;
  const int PS = fast ? F + 4 : F + 1, FP = PS;
  const int S = std::max(1, std::min(P, kThreads / F));                  // (filter, slab) items fill one round of the CTA
  const size_t inb = (static_cast<size_t>(T + 2) * (X + 2) * 2 + 3) & ~static_cast<size_t>(3);
  const size_t un_sparse = static_cast<size_t>(W) * C + 4u * C * F +
                           std::max(static_cast<size_t>(W) * 9 * F, static_cast<size_t>(S) * 18 * F);
  int dev = 0, max_smem = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  const size_t rest = inb + ((static_cast<size_t>(D) + 3) & ~static_cast<size_t>(3)) + ((18u * F + 3) & ~3u) +
                      ((static_cast<size_t>(P) * PS + 3) & ~static_cast<size_t>(3));
  const size_t tail = ((static_cast<size_t>(W) * C + 15) & ~static_cast<size_t>(15)) + 2 * static_cast<size_t>(P) * F;
  // the dense conv-1 gradient plane only where two CTAs per SM still fit
  const int dense_in = (rest + std::max(un_sparse, static_cast<size_t>(D) * FP) + 4) * sizeof(float) + tail <= 100u * 1024;
  const size_t un = dense_in ? std::max(un_sparse, static_cast<size_t>(D) * FP) : un_sparse;
  const size_t floats = inb + ((static_cast<size_t>(D) + 3) & ~static_cast<size_t>(3)) + ((18u * F + 3) & ~3u) + ((static_cast<size_t>(P) * PS + 3) & ~static_cast<size_t>(3)) +
                        ((un + 3) & ~static_cast<size_t>(3));
  const size_t smem = floats * sizeof(float) + ((static_cast<size_t>(W) * C + 15) & ~static_cast<size_t>(15)) +
                      2 * static_cast<size_t>(P) * F;
  if (T > 255 || X > 127) return -1;       // the packed location codes hold 8 + 7 bits
  if (smem > static_cast<size_t>(max_smem)) return -1;
  void (*kernel)(const TapeBwdParams) =
      !fast     ? conv3d_bwd_tape_kernel<0, 0>
      : F == 8  ? (X == 8 ? conv3d_bwd_tape_kernel<8, 8> : conv3d_bwd_tape_kernel<8, 0>)
      : F == 16 ? (X == 16 ? conv3d_bwd_tape_kernel<16, 16> : conv3d_bwd_tape_kernel<16, 0>)
                : (X == 32 ? conv3d_bwd_tape_kernel<32, 32> : conv3d_bwd_tape_kernel<32, 0>);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv3d_bwd_tape_kernel): %s", cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  TapeBwdParams p{a->w->w1, a->w->w2, F, a->in, a->mask, a->mask_invert, a->mask_alt, a->mask_invert_alt, a->row_set,
                  a->N, T, X, a->g_feat, a->g_in, a->partial, psize, a->tape->p1, a->tape->arg1, a->tape->arg2, S, dense_in};
  kernel<<<rows, kThreads, smem, stream>>>(p);
  return check_launch("conv3d_bwd_tape_kernel");
}

}  // namespace l2hmc
