// Backward pass of the ConvNet3D front-end (one input stream):
//   (N,T,X,2,1) -> Conv3D(F,(3,3,2),same,relu) -> MaxPool3D(2,2,same) -> Conv3D(2F,(2,2,2),same,relu)
//               -> MaxPool3D(2,2,same) -> flatten                       (reference l2hmc/network/conv_net.py:247-260)
// The reference gets these gradients from tf.gradients (gauge_model.py:823); here the adjoint of each layer is
// written out.  One chain per CTA pass, everything on-chip: the forward activations are RECOMPUTED from the chain's
// 2TX inputs (0.4 MFMA) instead of being kept on the tape (36 KB per chain and call at 16x16), then
//   pool-2 routing -> d conv-2 (weights, bias, input) -> pool-1 routing -> d conv-1 (weights, bias, input).
// Max-pool routes the gradient to the first maximum of its window; a maximum of 0 (all-ReLU-dead window) gets no
// gradient because relu'(0) = 0, so ties never matter.  Weight gradients accumulate into a per-CTA row of
// `partial` (chain n -> row n % rows, calls in stream order) and are summed over rows in fixed order by the caller:
// deterministic, no atomics.  TF 'same' semantics as in conv3d.cu (zero padding, odd cell at the end, clipped
// pooling windows; the kd = 1 taps of conv-2 only ever see padding, so their gradient is exactly zero).
#include "common.cuh"
#include "conv3d_bwd.cuh"

namespace l2hmc {
namespace {

constexpr int kThreads = 256;
constexpr int kMaxRows = 592;   // 4 resident CTAs per SM of the tape kernel (the recompute kernel fits 3: a short second wave)

struct ConvBwdParams {
  const float *w1, *b1, *w2, *b2;
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
};

// FT > 0: the filter count is a compile-time constant (index arithmetic folds, tap / channel loops unroll);
// FT = 0: any F.  XT > 0: the lattice extent X too (the reference builds its networks with F = space_size,
// gauge_dynamics.py:121-143, so X == F is the usual case): the (t, x, channel) decodes of every phase become shifts.
template <int FT, int XT>
__global__ void __launch_bounds__(kThreads, 3) conv3d_bwd_kernel(const ConvBwdParams p) {
  extern __shared__ __align__(16) float sm[];
  const int T = p.T, X = XT > 0 ? XT : p.X, F = FT > 0 ? FT : p.F, C2 = 2 * F;
  constexpr bool kFast = FT > 0 && FT % 8 == 0;   // register-tiled phases with 128-bit shared loads
  // padded row strides: lanes that walk rows hit distinct banks (+4 keeps float4 alignment on the fast path)
  const int FP = kFast ? F + 4 : F + 1, CP = kFast ? C2 + 4 : C2 + 1;
  const int T2 = (T + 1) / 2, X2 = (X + 1) / 2, T4 = (T2 + 1) / 2, X4 = (X2 + 1) / 2;
  const int D = 2 * T * X, K = T4 * X4 * C2;
  float* in_s = sm;                       // [T][X][2] (masked input)
  float* mk_s = in_s + D;                 // [T][X][2] the mask actually applied (1 if none)
  float* c1 = mk_s + D;                   // [T][X][2][F+1]    conv-1 output, later its gradient
  // pool-1 / conv-2 rows: +4 on the fast path too -- lanes that read 2 or 4 different positions with 128-bit loads
  // (rows 2F or 4F floats apart: the same banks) otherwise serialise
  const int PS = kFast ? F + 4 : F, CS = kFast ? C2 + 4 : C2;
  float* p1 = c1 + D * FP;                // [T2][X2][PS]      pool-1 output
  float* c2 = p1 + T2 * X2 * PS;          // [T2][X2][CS]      conv-2 output, later its gradient
  float* gp1 = c2 + T2 * X2 * CS;         // [T2][X2][F]       gradient of the pool-1 output
  float* w1s = gp1 + T2 * X2 * F;         // [3][3][2][F]
  float* w2s = w1s + 18 * F;              // [2][2][F][2F+1]   (kd = 0 taps)
  const int tid = threadIdx.x;
  for (int i = tid; i < 18 * F; i += kThreads) w1s[i] = p.w1[i];
  for (int i = tid; i < 4 * F * C2; i += kThreads) {
    const int co = i % C2, ci = (i / C2) % F, kk = i / (C2 * F);       // kk = kt*2 + kx
    w2s[(kk * F + ci) * CP + co] = p.w2[((kk * 2 + 0) * F + ci) * C2 + co];
  }
  float* part = p.partial + static_cast<size_t>(blockIdx.x) * p.psize;
  const int off_b1 = 18 * F, off_w2 = 19 * F, off_b2 = 19 * F + 8 * F * C2;

  for (int n = blockIdx.x; n < p.N; n += gridDim.x) {
    const bool alt = p.row_set != nullptr && p.row_set[n] != 0;
    const float* mask = alt ? p.mask_alt : p.mask;
    const int invert = alt ? p.invert_alt : p.invert;
    const float* src = p.in + static_cast<size_t>(n) * D;
    for (int i = tid; i < D; i += kThreads) {
      const float m = mask == nullptr ? 1.f : (invert ? 1.f - mask[i] : mask[i]);
      mk_s[i] = m;
      in_s[i] = m * src[i];
    }
    __syncthreads();
    // ---- forward, recomputed
    if (kFast && (X & 1) == 0) {       // thread = (two adjacent sites, 4 filters): a tap's weights feed 24 FMAs
      const int XH = X / 2;
      for (int i = tid; i < T * XH * (F / 4); i += kThreads) {
        const int f4 = (i % (F / 4)) * 4, pp = i / (F / 4), t = pp / XH, x = (pp - t * XH) * 2;
        float o0[4], o1[4], q0[4], q1[4];       // (site x: mu 0, mu 1), (site x + 1: mu 0, mu 1)
#pragma unroll
        for (int j = 0; j < 4; ++j) o0[j] = o1[j] = q0[j] = q1[j] = p.b1[f4 + j];
#pragma unroll
        for (int kt = 0; kt < 3; ++kt) {
          const int tt = t + kt - 1;
          if (tt < 0 || tt >= T) continue;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const int xa = x + kx - 1, xb = xa + 1;       // xa >= -1, xb <= X
            const float* row = in_s + (tt * X) * 2;
            const bool oka = xa >= 0, okb = xb < X;
            const float a0 = oka ? row[2 * xa] : 0.f, a1 = oka ? row[2 * xa + 1] : 0.f;
            const float b0 = okb ? row[2 * xb] : 0.f, b1 = okb ? row[2 * xb + 1] : 0.f;
            const float4 w0 = *reinterpret_cast<const float4*>(w1s + ((kt * 3 + kx) * 2) * F + f4);
            const float4 w1 = *reinterpret_cast<const float4*>(w1s + ((kt * 3 + kx) * 2 + 1) * F + f4);
            const float u0[4] = {w0.x, w0.y, w0.z, w0.w}, u1[4] = {w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              o0[j] = fmaf(a0, u0[j], o0[j]);       // mu = 0: taps kd = 0 (mu 0) and kd = 1 (mu 1)
              o0[j] = fmaf(a1, u1[j], o0[j]);
              o1[j] = fmaf(a1, u0[j], o1[j]);       // mu = 1: kd = 0 only (kd = 1 is padding)
              q0[j] = fmaf(b0, u0[j], q0[j]);
              q0[j] = fmaf(b1, u1[j], q0[j]);
              q1[j] = fmaf(b1, u0[j], q1[j]);
            }
          }
        }
        float* o = c1 + ((t * X + x) * 2) * FP + f4;
        *reinterpret_cast<float4*>(o) = make_float4(fmaxf(o0[0], 0.f), fmaxf(o0[1], 0.f), fmaxf(o0[2], 0.f), fmaxf(o0[3], 0.f));
        *reinterpret_cast<float4*>(o + FP) =
            make_float4(fmaxf(o1[0], 0.f), fmaxf(o1[1], 0.f), fmaxf(o1[2], 0.f), fmaxf(o1[3], 0.f));
        *reinterpret_cast<float4*>(o + 2 * FP) =
            make_float4(fmaxf(q0[0], 0.f), fmaxf(q0[1], 0.f), fmaxf(q0[2], 0.f), fmaxf(q0[3], 0.f));
        *reinterpret_cast<float4*>(o + 3 * FP) =
            make_float4(fmaxf(q1[0], 0.f), fmaxf(q1[1], 0.f), fmaxf(q1[2], 0.f), fmaxf(q1[3], 0.f));
      }
    } else if (kFast) {       // thread = (site, 4 filters): both mu outputs, weights as float4
      for (int i = tid; i < T * X * (F / 4); i += kThreads) {
        const int f4 = (i % (F / 4)) * 4, site = i / (F / 4), t = site / X, x = site - t * X;
        float o0[4], o1[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o0[j] = o1[j] = p.b1[f4 + j];
#pragma unroll
        for (int kt = 0; kt < 3; ++kt) {
          const int tt = t + kt - 1;
          if (tt < 0 || tt >= T) continue;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const int xx = x + kx - 1;
            if (xx < 0 || xx >= X) continue;
            const float in0 = in_s[(tt * X + xx) * 2], in1 = in_s[(tt * X + xx) * 2 + 1];
            const float4 w0 = *reinterpret_cast<const float4*>(w1s + ((kt * 3 + kx) * 2) * F + f4);
            const float4 w1 = *reinterpret_cast<const float4*>(w1s + ((kt * 3 + kx) * 2 + 1) * F + f4);
            const float a0[4] = {w0.x, w0.y, w0.z, w0.w}, a1[4] = {w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              o0[j] = fmaf(in0, a0[j], o0[j]);       // mu = 0: taps kd = 0 (mu 0) and kd = 1 (mu 1)
              o0[j] = fmaf(in1, a1[j], o0[j]);
              o1[j] = fmaf(in1, a0[j], o1[j]);       // mu = 1: kd = 0 only (kd = 1 is padding)
            }
          }
        }
        *reinterpret_cast<float4*>(c1 + (site * 2) * FP + f4) =
            make_float4(fmaxf(o0[0], 0.f), fmaxf(o0[1], 0.f), fmaxf(o0[2], 0.f), fmaxf(o0[3], 0.f));
        *reinterpret_cast<float4*>(c1 + (site * 2 + 1) * FP + f4) =
            make_float4(fmaxf(o1[0], 0.f), fmaxf(o1[1], 0.f), fmaxf(o1[2], 0.f), fmaxf(o1[3], 0.f));
      }
    } else
    for (int i = tid; i < D * F; i += kThreads) {
      const int f = i % F, pos = i / F, d = pos & 1, site = pos >> 1, t = site / X, x = site - t * X;
      float acc = p.b1[f];
#pragma unroll
      for (int kt = 0; kt < 3; ++kt) {
        const int tt = t + kt - 1;
        if (tt < 0 || tt >= T) continue;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int xx = x + kx - 1;
          if (xx < 0 || xx >= X) continue;
          const float* src2 = in_s + (tt * X + xx) * 2;
          const float* wk = w1s + ((kt * 3 + kx) * 2) * F + f;
          acc = fmaf(src2[d], wk[0], acc);                 // kd = 0
          if (d == 0) acc = fmaf(src2[1], wk[F], acc);     // kd = 1 (mu = 1 sees only padding there)
        }
      }
      c1[pos * FP + f] = fmaxf(acc, 0.f);
    }
    __syncthreads();
    for (int i = tid; i < T2 * X2 * F; i += kThreads) {
      const int f = i % F, x2 = (i / F) % X2, t2 = i / (F * X2);
      float best = -INFINITY;
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
          const int t = 2 * t2 + a, x = 2 * x2 + b;
          if (t >= T || x >= X) continue;
          for (int d = 0; d < 2; ++d) best = fmaxf(best, c1[((t * X + x) * 2 + d) * FP + f]);
        }
      p1[(t2 * X2 + x2) * PS + f] = best;
    }
    __syncthreads();
    if (kFast && (X2 & 1) == 0) {       // thread = (two adjacent positions, 4 output channels): one weight load, 8 FMAs
      const int XH = X2 / 2;
      for (int i = tid; i < T2 * XH * (C2 / 4); i += kThreads) {
        const int co4 = (i % (C2 / 4)) * 4, pp = i / (C2 / 4), x = (pp % XH) * 2, t = pp / XH;
        float acc[4], bcc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = bcc[j] = p.b2[co4 + j];
        for (int kt = 0; kt < 2; ++kt)
          for (int kx = 0; kx < 2; ++kx) {
            const int tt = t + kt, xx = x + kx;          // xx <= X2 - 1; the second position's tap may fall outside
            if (tt >= T2) continue;
            const bool okb = xx + 1 < X2;
            const float* a = p1 + (tt * X2 + xx) * PS;
            const float* w = w2s + ((kt * 2 + kx) * F) * CP + co4;
#pragma unroll
            for (int c4 = 0; c4 < F; c4 += 4) {
              const float4 av = *reinterpret_cast<const float4*>(a + c4);
              const float4 bv = okb ? *reinterpret_cast<const float4*>(a + PS + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
              const float aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float4 wv = *reinterpret_cast<const float4*>(w + (c4 + q) * CP);
                acc[0] = fmaf(aa[q], wv.x, acc[0]), acc[1] = fmaf(aa[q], wv.y, acc[1]);
                acc[2] = fmaf(aa[q], wv.z, acc[2]), acc[3] = fmaf(aa[q], wv.w, acc[3]);
                bcc[0] = fmaf(bb[q], wv.x, bcc[0]), bcc[1] = fmaf(bb[q], wv.y, bcc[1]);
                bcc[2] = fmaf(bb[q], wv.z, bcc[2]), bcc[3] = fmaf(bb[q], wv.w, bcc[3]);
              }
            }
          }
        float* o = c2 + (t * X2 + x) * CS + co4;
        *reinterpret_cast<float4*>(o) =
            make_float4(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
        *reinterpret_cast<float4*>(o + CS) =
            make_float4(fmaxf(bcc[0], 0.f), fmaxf(bcc[1], 0.f), fmaxf(bcc[2], 0.f), fmaxf(bcc[3], 0.f));
      }
    } else if (kFast) {       // thread = (position, 4 output channels)
      for (int i = tid; i < T2 * X2 * (C2 / 4); i += kThreads) {
        const int co4 = (i % (C2 / 4)) * 4, pos = i / (C2 / 4), x = pos % X2, t = pos / X2;
        float acc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = p.b2[co4 + j];
        for (int kt = 0; kt < 2; ++kt)
          for (int kx = 0; kx < 2; ++kx) {
            const int tt = t + kt, xx = x + kx;
            if (tt >= T2 || xx >= X2) continue;
            const float* a = p1 + (tt * X2 + xx) * PS;
            const float* w = w2s + ((kt * 2 + kx) * F) * CP + co4;
#pragma unroll
            for (int c4 = 0; c4 < F; c4 += 4) {
              const float4 av = *reinterpret_cast<const float4*>(a + c4);
              const float aa[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float4 wv = *reinterpret_cast<const float4*>(w + (c4 + q) * CP);
                acc[0] = fmaf(aa[q], wv.x, acc[0]), acc[1] = fmaf(aa[q], wv.y, acc[1]);
                acc[2] = fmaf(aa[q], wv.z, acc[2]), acc[3] = fmaf(aa[q], wv.w, acc[3]);
              }
            }
          }
        *reinterpret_cast<float4*>(c2 + pos * CS + co4) =
            make_float4(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
      }
    } else
    for (int i = tid; i < T2 * X2 * C2; i += kThreads) {
      const int co = i % C2, x = (i / C2) % X2, t = i / (C2 * X2);
      float acc = p.b2[co];
      for (int kt = 0; kt < 2; ++kt)
        for (int kx = 0; kx < 2; ++kx) {
          const int tt = t + kt, xx = x + kx;
          if (tt >= T2 || xx >= X2) continue;
          const float* a = p1 + (tt * X2 + xx) * PS;
          const float* w = w2s + ((kt * 2 + kx) * F) * CP + co;
#pragma unroll 8
          for (int ci = 0; ci < F; ++ci) acc = fmaf(a[ci], w[ci * CP], acc);
        }
      c2[i] = fmaxf(acc, 0.f);
    }
    __syncthreads();
    // ---- pool-2 routing: c2 <- gradient of the conv-2 pre-activation
    const float* gf = p.g_feat + static_cast<size_t>(n) * K;
    for (int i = tid; i < T4 * X4 * C2; i += kThreads) {
      const int co = i % C2, x4 = (i / C2) % X4, t4 = i / (C2 * X4);
      float best = -INFINITY;
      int arg = -1;
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
          const int t = 2 * t4 + a, x = 2 * x4 + b;
          if (t >= T2 || x >= X2) continue;
          const int at = (t * X2 + x) * CS + co;
          if (c2[at] > best) best = c2[at], arg = at;
        }
      const float g = best > 0.f ? gf[i] : 0.f;
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
          const int t = 2 * t4 + a, x = 2 * x4 + b;
          if (t >= T2 || x >= X2) continue;
          const int at = (t * X2 + x) * CS + co;
          c2[at] = at == arg ? g : 0.f;
        }
    }
    __syncthreads();
    // ---- conv-2: weight / bias gradients (kd = 0 taps), input gradient
    if (kFast) {       // thread = (tap, output channel, 8 input channels): one gradient load feeds 8 FMAs
      for (int i = tid; i < 4 * C2 * (F / 8); i += kThreads) {
        const int co = i % C2, ci0 = ((i / C2) % (F / 8)) * 8, kk = i / (C2 * (F / 8)), kt = kk >> 1, kx = kk & 1;
        float s8[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) s8[j] = 0.f;
        for (int t = 0; t + kt < T2; ++t) {
          const float* a = p1 + ((t + kt) * X2 + kx) * PS + ci0;
          const float* g = c2 + (t * X2) * CS + co;
          const int nx = X2 - kx;
#pragma unroll 2
          for (int x = 0; x < nx; ++x) {
            const float gv = g[x * CS];
            const float4 a0 = *reinterpret_cast<const float4*>(a + x * PS);
            const float4 a1 = *reinterpret_cast<const float4*>(a + x * PS + 4);
            s8[0] = fmaf(a0.x, gv, s8[0]), s8[1] = fmaf(a0.y, gv, s8[1]), s8[2] = fmaf(a0.z, gv, s8[2]);
            s8[3] = fmaf(a0.w, gv, s8[3]), s8[4] = fmaf(a1.x, gv, s8[4]), s8[5] = fmaf(a1.y, gv, s8[5]);
            s8[6] = fmaf(a1.z, gv, s8[6]), s8[7] = fmaf(a1.w, gv, s8[7]);
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) part[off_w2 + ((kk * 2 + 0) * F + ci0 + j) * C2 + co] += s8[j];
      }
    } else
    for (int i = tid; i < 4 * F * C2; i += kThreads) {
      const int co = i % C2, ci = (i / C2) % F, kk = i / (C2 * F), kt = kk >> 1, kx = kk & 1;
      float s = 0.f;
      for (int t = 0; t + kt < T2; ++t) {
        const float* a = p1 + ((t + kt) * X2 + kx) * PS + ci;
        const float* g = c2 + (t * X2) * CS + co;
        const int nx = X2 - kx;
#pragma unroll 4
        for (int x = 0; x < nx; ++x) s = fmaf(a[x * PS], g[x * CS], s);
      }
      part[off_w2 + ((kk * 2 + 0) * F + ci) * C2 + co] += s;
    }
    for (int co = tid; co < C2; co += kThreads) {
      float s = 0.f;
      for (int j = 0; j < T2 * X2; ++j) s += c2[j * CS + co];
      part[off_b2 + co] += s;
    }
    if (kFast && (X2 & 1) == 0) {
      // thread = (two adjacent positions, input channel): the weight row of a tap is loaded once for both
      const int XH = X2 / 2;
      for (int i = tid; i < T2 * XH * F; i += kThreads) {
        const int ci = i % F, x1 = ((i / F) % XH) * 2 + 1, t = i / (F * XH);      // positions x1 - 1 and x1
        float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};
        for (int kt = 0; kt < 2 && kt <= t; ++kt)
          for (int kx = 0; kx < 2; ++kx) {
            const bool oka = x1 - 1 - kx >= 0;
            const float* gb = c2 + ((t - kt) * X2 + x1 - kx) * CS;          // gradient row read by position x1
            const float* w = w2s + ((kt * 2 + kx) * F + ci) * CP;
#pragma unroll
            for (int c4 = 0; c4 < C2; c4 += 4) {
              const float4 wv = *reinterpret_cast<const float4*>(w + c4);
              const float4 gv = *reinterpret_cast<const float4*>(gb + c4);
              const float4 ga = oka ? *reinterpret_cast<const float4*>(gb - CS + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
              sb[0] = fmaf(wv.x, gv.x, sb[0]), sb[1] = fmaf(wv.y, gv.y, sb[1]);
              sb[2] = fmaf(wv.z, gv.z, sb[2]), sb[3] = fmaf(wv.w, gv.w, sb[3]);
              sa[0] = fmaf(wv.x, ga.x, sa[0]), sa[1] = fmaf(wv.y, ga.y, sa[1]);
              sa[2] = fmaf(wv.z, ga.z, sa[2]), sa[3] = fmaf(wv.w, ga.w, sa[3]);
            }
          }
        float* o = gp1 + (t * X2 + x1 - 1) * F + ci;
        o[0] = (sa[0] + sa[1]) + (sa[2] + sa[3]);
        o[F] = (sb[0] + sb[1]) + (sb[2] + sb[3]);
      }
    } else
    for (int i = tid; i < T2 * X2 * F; i += kThreads) {
      const int ci = i % F, x = (i / F) % X2, t = i / (F * X2);
      float s = 0.f;
      float s4[4] = {0.f, 0.f, 0.f, 0.f};       // fast path: four independent chains, joined in fixed order below
      for (int kt = 0; kt < 2 && kt <= t; ++kt)
        for (int kx = 0; kx < 2 && kx <= x; ++kx) {
          const float* g = c2 + ((t - kt) * X2 + x - kx) * CS;
          const float* w = w2s + ((kt * 2 + kx) * F + ci) * CP;
          if (kFast) {
#pragma unroll
            for (int c4 = 0; c4 < C2; c4 += 4) {
              const float4 wv = *reinterpret_cast<const float4*>(w + c4);
              const float4 gv = *reinterpret_cast<const float4*>(g + c4);
              s4[0] = fmaf(wv.x, gv.x, s4[0]), s4[1] = fmaf(wv.y, gv.y, s4[1]);
              s4[2] = fmaf(wv.z, gv.z, s4[2]), s4[3] = fmaf(wv.w, gv.w, s4[3]);
            }
          } else {
#pragma unroll 8
            for (int co = 0; co < C2; ++co) s = fmaf(w[co], g[co], s);
          }
        }
      if (kFast) s = (s4[0] + s4[1]) + (s4[2] + s4[3]);
      gp1[i] = s;
    }
    __syncthreads();
    // ---- pool-1 routing: c1 <- gradient of the conv-1 pre-activation
    for (int i = tid; i < T2 * X2 * F; i += kThreads) {
      const int f = i % F, x2 = (i / F) % X2, t2 = i / (F * X2);
      float best = -INFINITY;
      int arg = -1;
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
          const int t = 2 * t2 + a, x = 2 * x2 + b;
          if (t >= T || x >= X) continue;
          for (int d = 0; d < 2; ++d) {
            const int at = ((t * X + x) * 2 + d) * FP + f;
            if (c1[at] > best) best = c1[at], arg = at;
          }
        }
      const float g = best > 0.f ? gp1[i] : 0.f;
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
          const int t = 2 * t2 + a, x = 2 * x2 + b;
          if (t >= T || x >= X) continue;
          for (int d = 0; d < 2; ++d) {
            const int at = ((t * X + x) * 2 + d) * FP + f;
            c1[at] = at == arg ? g : 0.f;
          }
        }
    }
    __syncthreads();
    // ---- conv-1: weight / bias gradients, input gradient
    if (kFast) {
      // thread = (row slab, tap, 4 filters): both kd entries from one pass over the slab's outputs; the slabs'
      // partial sums meet in the (now free) conv-2 buffers and are added in slab order
      int S = 1;
      while (S * 2 <= 8 && T % (S * 2) == 0 && (S * 2) * 18 * F <= T2 * X2 * (C2 + F)) S *= 2;
      float* scratch = c2;
      const int rows_per = T / S;
      for (int i = tid; i < S * 9 * (F / 4); i += kThreads) {
        const int f4 = (i % (F / 4)) * 4, tap = (i / (F / 4)) % 9, slab = i / ((F / 4) * 9), kt = tap / 3, kx = tap - 3 * kt;
        float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
        const int t_lo = max(kt == 0 ? 1 : 0, slab * rows_per), t_hi = min(kt == 2 ? T - 1 : T, (slab + 1) * rows_per);
        const int x_lo = kx == 0 ? 1 : 0, x_hi = kx == 2 ? X - 1 : X;
        for (int t = t_lo; t < t_hi; ++t) {
          const float* a = in_s + ((t + kt - 1) * X + kx - 1) * 2;
          const float* g = c1 + (t * X) * 2 * FP + f4;
#pragma unroll 2
          for (int x = x_lo; x < x_hi; ++x) {
            const float in0 = a[2 * x], in1 = a[2 * x + 1];
            const float4 g0 = *reinterpret_cast<const float4*>(g + (2 * x) * FP);
            const float4 g1 = *reinterpret_cast<const float4*>(g + (2 * x + 1) * FP);
            const float b0[4] = {g0.x, g0.y, g0.z, g0.w}, b1[4] = {g1.x, g1.y, g1.z, g1.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              s0[j] = fmaf(in0, b0[j], s0[j]);       // kd = 0: output mu = 0 reads input mu = 0, mu = 1 reads mu = 1
              s0[j] = fmaf(in1, b1[j], s0[j]);
              s1[j] = fmaf(in1, b0[j], s1[j]);       // kd = 1: output mu = 0 reads input mu = 1
            }
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          scratch[slab * 18 * F + (tap * 2) * F + f4 + j] = s0[j];
          scratch[slab * 18 * F + (tap * 2 + 1) * F + f4 + j] = s1[j];
        }
      }
      // bias gradient: thread = (slab of outputs, filter), taken from the top of the CTA (the threads the last pass
      // above leaves idle); the slab sums meet in the (now free) pool-1 buffer
      const int bs = min(kThreads / F, min(8, T2 * X2)), b_per = (D + bs - 1) / bs;
      {
        const int j = kThreads - 1 - tid;
        if (j < bs * F) {
          const int f = j % F, slab = j / F;
          float sb = 0.f;
          for (int r = slab * b_per; r < min(D, (slab + 1) * b_per); ++r) sb += c1[r * FP + f];
          p1[j] = sb;
        }
      }
      __syncthreads();
      for (int i = tid; i < 18 * F; i += kThreads) {
        float s = 0.f;
        for (int q = 0; q < S; ++q) s += scratch[q * 18 * F + i];
        part[i] += s;
      }
      for (int f = tid; f < F; f += kThreads) {
        float s = 0.f;
        for (int q = 0; q < bs; ++q) s += p1[q * F + f];
        part[off_b1 + f] += s;
      }
    } else
    for (int i = tid; i < 18 * F; i += kThreads) {
      const int f = i % F, kd = (i / F) & 1, kx = (i / (2 * F)) % 3, kt = i / (6 * F);
      float s = 0.f;
      const int t_lo = kt == 0 ? 1 : 0, t_hi = kt == 2 ? T - 1 : T;       // outputs whose tap lands inside the lattice
      const int x_lo = kx == 0 ? 1 : 0, x_hi = kx == 2 ? X - 1 : X;
      for (int t = t_lo; t < t_hi; ++t) {
        const float* a = in_s + ((t + kt - 1) * X + kx - 1) * 2 + kd;
        const float* g = c1 + (t * X) * 2 * FP + f;
        if (kd == 0) {
#pragma unroll 4
          for (int x = x_lo; x < x_hi; ++x) {
            s = fmaf(a[2 * x], g[(2 * x) * FP], s);
            s = fmaf(a[2 * x + 1], g[(2 * x + 1) * FP], s);
          }
        } else {
#pragma unroll 4
          for (int x = x_lo; x < x_hi; ++x) s = fmaf(a[2 * x], g[(2 * x) * FP], s);
        }
      }
      part[i] += s;
    }
    if (!kFast)
      for (int f = tid; f < F; f += kThreads) {
        float s = 0.f;
        for (int j = 0; j < D; ++j) s += c1[j * FP + f];
        part[off_b1 + f] += s;
      }
    float* dst = p.g_in + static_cast<size_t>(n) * D;
    if (kFast) {
      // thread = input site (t, x), both mu: mu = 0 gets tap kd = 0 of output mu = 0; mu = 1 gets kd = 0 of output
      // mu = 1 and kd = 1 of output mu = 0 -- the two output rows and the two weight rows are loaded once per tap
      for (int site = tid; site < T * X; site += kThreads) {
        const int xx = site % X, tt = site / X;
        float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
        for (int kt = 0; kt < 3; ++kt) {
          const int t = tt - kt + 1;
          if (t < 0 || t >= T) continue;
          for (int kx = 0; kx < 3; ++kx) {
            const int x = xx - kx + 1;
            if (x < 0 || x >= X) continue;
            const float* g = c1 + ((t * X + x) * 2) * FP;
            const float* w = w1s + ((kt * 3 + kx) * 2) * F;
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
          }
        }
        const int i0 = 2 * site;
        dst[i0] += mk_s[i0] * ((s0[0] + s0[1]) + (s0[2] + s0[3]));
        dst[i0 + 1] += mk_s[i0 + 1] * ((s1[0] + s1[1]) + (s1[2] + s1[3]));
      }
    } else
    for (int i = tid; i < D; i += kThreads) {
      const int dd = i & 1, xx = (i >> 1) % X, tt = i / (2 * X);     // input position
      float s = 0.f;
      for (int kt = 0; kt < 3; ++kt) {
        const int t = tt - kt + 1;
        if (t < 0 || t >= T) continue;
        for (int kx = 0; kx < 3; ++kx) {
          const int x = xx - kx + 1;
          if (x < 0 || x >= X) continue;
          for (int kd = 0; kd <= dd; ++kd) {
            const float* g = c1 + ((t * X + x) * 2 + dd - kd) * FP;
            const float* w = w1s + ((kt * 3 + kx) * 2 + kd) * F;
            if (kFast) {
#pragma unroll
              for (int f4 = 0; f4 < F; f4 += 4) {
                const float4 wv = *reinterpret_cast<const float4*>(w + f4);
                const float4 gv = *reinterpret_cast<const float4*>(g + f4);
                s = fmaf(wv.x, gv.x, s), s = fmaf(wv.y, gv.y, s), s = fmaf(wv.z, gv.z, s), s = fmaf(wv.w, gv.w, s);
              }
            } else {
#pragma unroll 8
              for (int f = 0; f < F; ++f) s = fmaf(w[f], g[f], s);
            }
          }
        }
      }
      dst[i] += mk_s[i] * s;
    }
    __syncthreads();
  }
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_conv3d_bwd_rows(int N) { return N < kMaxRows ? (N > 0 ? N : 1) : kMaxRows; }

extern "C" int l2hmc_conv3d_bwd_cols(int F) { return 18 * F + F + 8 * F * 2 * F + 2 * F; }

extern "C" int l2hmc_conv3d_backward(const l2hmc_conv3d_bwd_args* a, void* stream) {
  L2HMC_REQUIRE(a && a->w, "l2hmc_conv3d_backward: null arguments");
  if (a->N == 0) return L2HMC_OK;
  const int F = a->w->F, T = a->T, X = a->X;
  L2HMC_REQUIRE(a->N > 0 && T > 0 && X > 0 && F > 0, "l2hmc_conv3d_backward: bad sizes N=%d T=%d X=%d F=%d", a->N, T, X, F);
  L2HMC_REQUIRE(a->w->w1 && a->w->b1 && a->w->w2 && a->w->b2 && a->in && a->g_feat && a->g_in && a->partial,
                "l2hmc_conv3d_backward: null pointer");
  L2HMC_REQUIRE(!a->row_set || (!a->mask == !a->mask_alt), "l2hmc_conv3d_backward: mask and mask_alt go together");
  if (a->tape != nullptr) {      // the forward pass's routing tape: nothing recomputed, products over the non-zeros only
    L2HMC_REQUIRE(a->tape->p1 && a->tape->arg1 && a->tape->arg2, "l2hmc_conv3d_backward: null tape array");
    const int rc = conv3d_backward_tape(a, l2hmc_conv3d_bwd_rows(a->N), l2hmc_conv3d_bwd_cols(F),
                                        static_cast<cudaStream_t>(stream));
    if (rc >= 0) return rc;
    set_error("l2hmc_conv3d_backward: %dx%d with F=%d does not fit shared memory even with the tape", T, X, F);
    return L2HMC_ERR_UNSUPPORTED;
  }
  const int T2 = (T + 1) / 2, X2 = (X + 1) / 2, D = 2 * T * X;
  const bool fast = (F == 8 || F == 16 || F == 32) && D % 4 == 0;   // float4 rows need 16-byte aligned buffers
  const size_t fp = fast ? F + 4 : F + 1, cp = fast ? 2 * F + 4 : 2 * F + 1;
  const size_t ps = fast ? F + 4 : F, cs = fast ? 2 * F + 4 : 2 * F;       // pool-1 / conv-2 row strides
  const size_t floats = 2u * D + static_cast<size_t>(D) * fp + static_cast<size_t>(T2) * X2 * (ps + F + cs) + 18u * F +
                        4u * F * cp;
  const size_t smem = floats * sizeof(float);
  int dev = 0, max_smem = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (smem > static_cast<size_t>(max_smem)) {
    set_error("l2hmc_conv3d_backward: %dx%d with F=%d needs %zu B of shared memory per chain (limit %d B)", T, X, F, smem,
              max_smem);
    return L2HMC_ERR_UNSUPPORTED;
  }
  void (*kernel)(const ConvBwdParams) =
      !fast     ? conv3d_bwd_kernel<0, 0>
      : F == 8  ? (X == 8 ? conv3d_bwd_kernel<8, 8> : conv3d_bwd_kernel<8, 0>)
      : F == 16 ? (X == 16 ? conv3d_bwd_kernel<16, 16> : conv3d_bwd_kernel<16, 0>)
                : (X == 32 ? conv3d_bwd_kernel<32, 32> : conv3d_bwd_kernel<32, 0>);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv3d_bwd_kernel): %s", cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  ConvBwdParams p{a->w->w1, a->w->b1, a->w->w2, a->w->b2, F, a->in, a->mask, a->mask_invert, a->mask_alt,
                  a->mask_invert_alt, a->row_set, a->N, T, X, a->g_feat, a->g_in, a->partial, l2hmc_conv3d_bwd_cols(F)};
  kernel<<<l2hmc_conv3d_bwd_rows(a->N), kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("conv3d_bwd_kernel");
}
