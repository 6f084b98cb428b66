// the backward pass of the training loss and the optimiser step.
//
// Replaces (reference l2hmc/...), all of which TensorFlow derives by graph autodiff (`tf.gradients`,
// gauge_model.py:805-832) -- here every adjoint is written out:
//   gauge_model.py:762-801                 d loss / d (x_prop, px, pz)            loss_bwd_kernel
//   dynamics/gauge_dynamics.py:592-609     accept probability                     accept_bwd_kernel, traj_bwd_init_kernel
//   dynamics/gauge_dynamics.py:486-590     the four augmented-leapfrog updates    momentum_bwd_kernel / position_bwd_kernel
//   dynamics/gauge_dynamics.py:698-709     force = tf.gradients(beta*S) -> its    force_hvp_kernel (Hessian-vector product,
//                                          own adjoint is a second derivative     same plaquette stencil applied twice)
//   network/generic_net.py:129-146         dense trunk + heads                    sgemm_kernel (NT / TN forms, ReLU gate,
//                                                                                 column-mask and accumulate epilogues)
//   gauge_model.py:925-945                 tf.train.AdamOptimizer [+ :826-827     adam_kernel (one launch over the flat
//                                          clip_by_global_norm]                   parameter arena)
//
// Design (B200): the forward pass of a training step records a TAPE in HBM (per network call: both inputs, h1, h2,
// S, T, Q; x- and z-chain stacked into one batch of 2N rows, forward and backward chains in the same launches via
// row_set).  The backward sweep produces only the per-call activation gradients (dA, dZ2, dZ1) and appends them to
// the tape; every WEIGHT gradient is one deferred TN GEMM over the whole tape (K = calls x rows instead of `calls`
// small GEMMs), split along K with a fixed-order second stage, so results are deterministic (no atomics anywhere).
// The small-batch products run on the FMA pipe (sgemm_kernel here for the strided forms, the chunked dense kernel of
// nets.cu for d act = d out . W^T); dense contractions go to tcgen05 (gemm_tc.cu, dispatch l2hmc_gemm).
#include <cooperative_groups.h>

#include <algorithm>
#include <atomic>

#include "common.cuh"
#include "nets.cuh"
#include "update_math.cuh"

namespace l2hmc {
namespace {

// =============================================================================================== SGEMM
constexpr int GBM = 64, GBN = 64, GBK = 16, kGemmThreads = 256;

struct GemmParams {
  l2hmc_gemm_args a;
  int ksplit;
  float* partial;   // [ksplit][M*N] when ksplit > 1 and the K ranges are not a cluster
  int cluster;      // 1: the ksplit (<= 8) CTAs of a tile form a thread-block cluster along z; rank 0 adds the other
                    // ranks' accumulators out of their shared memory in rank order (the same order as
                    // splitk_reduce_kernel) and runs the epilogue -- no partial round trip, no second launch
};

// 4-byte async copy global -> shared (LDGSTS), zero-filled when !valid: operand tiles are transposed on the fly
// (any strides), so the copy granularity is one float.
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc, bool valid) {
  const int sz = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int kStages = 5;   // K slabs in flight: the training GEMMs at config 2 are 32-CTA launches, latency-bound

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}

// A_KC / B_KC: the operand's K index is the contiguous one (only affects which threads read neighbouring words).
// VEC4: A^T . B with A [K, M] and B [K, N] row-major, 16-byte aligned rows and M, N multiples of 4 (the weight
// gradients over the tape): one 16-byte async copy per thread, operand and slab instead of four 4-byte ones.
template <bool A_KC, bool B_KC, bool VEC4 = false>
__global__ void __launch_bounds__(kGemmThreads) sgemm_kernel(const GemmParams p) {
  pdl_wait();
  pdl_trigger();
  __shared__ __align__(16) float As[kStages][GBK][GBM + 4];
  __shared__ __align__(16) float Bs[kStages][GBK][GBN + 4];
  const l2hmc_gemm_args& g = p.a;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * GBM, n0 = blockIdx.x * GBN;
  const int z = blockIdx.z;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  int am[4], ak[4], bk[4], bn[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int e = tid + i * kGemmThreads;
    if (A_KC) ak[i] = e & 15, am[i] = e >> 4; else am[i] = e & 63, ak[i] = e >> 6;
    if (B_KC) bk[i] = e & 15, bn[i] = e >> 4; else bn[i] = e & 63, bk[i] = e >> 6;
  }

  // K range of segment s for this CTA (split-K slices segment 0 over grid.z)
  auto seg_range = [&](int s, int& kb, int& ke) {
    kb = 0, ke = g.seg[s].K;
    if (p.ksplit > 1) {
      int chunk = (g.seg[s].K + p.ksplit - 1) / p.ksplit;
      chunk = (chunk + GBK - 1) / GBK * GBK;
      kb = min(g.seg[s].K, z * chunk);
      ke = min(g.seg[s].K, kb + chunk);
    }
  };
  int total = 0;
  for (int s = 0; s < g.nseg; ++s) {
    int kb, ke;
    seg_range(s, kb, ke);
    total += (ke - kb + GBK - 1) / GBK;
  }
  // issue cursor over the flattened (segment, slab) sequence
  int iseg = 0, ik0 = 0, ike = 0;
  seg_range(0, ik0, ike);
  auto skip_empty = [&]() {
    while (iseg < g.nseg && ik0 >= ike) {
      ++iseg;
      if (iseg < g.nseg) seg_range(iseg, ik0, ike);
    }
  };
  skip_empty();
  auto issue = [&](int stage) {
    if (iseg < g.nseg && VEC4) {
      const l2hmc_gemm_seg& sg = g.seg[iseg];
      const int kk = tid >> 4, c4 = (tid & 15) * 4, k = ik0 + kk;          // 16 K rows x 16 float4 per operand
      const bool va = k < ike && m0 + c4 < g.M, vb = k < ike && n0 + c4 < g.N;
      cp_async16(&As[stage][kk][c4], va ? sg.A + static_cast<long long>(k) * sg.a_cs + (m0 + c4) : sg.A, va);
      cp_async16(&Bs[stage][kk][c4], vb ? sg.B + static_cast<long long>(k) * sg.b_rs + (n0 + c4) : sg.B, vb);
      ik0 += GBK;
      skip_empty();
    } else if (iseg < g.nseg) {
      const l2hmc_gemm_seg& sg = g.seg[iseg];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = m0 + am[i], k = ik0 + ak[i];
        const bool va = m < g.M && k < ike;
        cp_async4(&As[stage][ak[i]][am[i]],
                  va ? sg.A + static_cast<long long>(m) * sg.a_rs + static_cast<long long>(k) * sg.a_cs : sg.A, va);
        const int n = n0 + bn[i], k2 = ik0 + bk[i];
        const bool vb = n < g.N && k2 < ike;
        cp_async4(&Bs[stage][bk[i]][bn[i]],
                  vb ? sg.B + static_cast<long long>(k2) * sg.b_rs + static_cast<long long>(n) * sg.b_cs : sg.B, vb);
      }
      ik0 += GBK;
      skip_empty();
    }
    cp_async_commit();
  };
#pragma unroll
  for (int s = 0; s < kStages - 1; ++s) issue(s);
  for (int it = 0; it < total; ++it) {
    cp_async_wait<kStages - 2>();
    __syncthreads();                       // slab `it` has landed; everyone is done with slab it - 1
    issue((it + kStages - 1) % kStages);   // refill the stage slab it - 1 occupied
    const int st = it % kStages;
#pragma unroll
    for (int k = 0; k < GBK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[st][k][4 * ty]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[st][k][4 * tx]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }
  cp_async_wait<0>();

  if (p.cluster) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    float* red = &As[0][0][0];             // 64 x 64 accumulators; the operand ring is drained
    __syncthreads();
    if (z != 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(red + (4 * ty + i) * GBN + 4 * tx) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    cl.sync();
    if (z == 0) {
      for (int q = 1; q < p.ksplit; ++q) {
        const float* remote = cl.map_shared_rank(red, q);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 v = *reinterpret_cast<const float4*>(remote + (4 * ty + i) * GBN + 4 * tx);
          acc[i][0] = __fadd_rn(acc[i][0], v.x), acc[i][1] = __fadd_rn(acc[i][1], v.y);
          acc[i][2] = __fadd_rn(acc[i][2], v.z), acc[i][3] = __fadd_rn(acc[i][3], v.w);
        }
      }
    }
    cl.sync();                             // the other ranks' shared memory stays alive until rank 0 has read it
    if (z != 0) return;
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = m0 + 4 * ty + i;
    if (r >= g.M) continue;
    const bool alt = g.row_set != nullptr && g.row_set[r] != 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = n0 + 4 * tx + j;
      if (c >= g.N) continue;
      float val = acc[i][j];
      if (p.ksplit > 1 && !p.cluster) {
        p.partial[(static_cast<size_t>(z) * g.M + r) * g.N + c] = val;
        continue;
      }
      if (g.relu_gate != nullptr && !(g.relu_gate[static_cast<long long>(r) * g.ldg + c] > 0.f)) val = 0.f;
      const float* cs = alt ? g.col_scale_alt : g.col_scale;
      if (cs != nullptr) {
        const float sc = cs[c];
        val = __fmul_rn(val, (alt ? g.col_scale_invert_alt : g.col_scale_invert) ? __fsub_rn(1.f, sc) : sc);
      }
      float* dst = g.C + static_cast<long long>(r) * g.ldc + c;
      if (g.accumulate) val = __fadd_rn(*dst, val);
      *dst = val;
    }
  }
}

// C[r][c] (=|+=) scale[c] * gate[r][c] * sum_z partial[z][r][c], z in fixed order (the epilogue of a split-K product)
__global__ void splitk_reduce_kernel(const float* __restrict__ partial, int ksplit, const l2hmc_gemm_args g) {
  pdl_wait();
  pdl_trigger();
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t total = static_cast<size_t>(g.M) * g.N;
  if (idx >= total) return;
  float val = 0.f;
  for (int z = 0; z < ksplit; ++z) val = __fadd_rn(val, partial[static_cast<size_t>(z) * total + idx]);
  const int r = static_cast<int>(idx / g.N), c = static_cast<int>(idx % g.N);
  if (g.relu_gate != nullptr && !(g.relu_gate[static_cast<long long>(r) * g.ldg + c] > 0.f)) val = 0.f;
  const bool alt = g.row_set != nullptr && g.row_set[r] != 0;
  const float* cs = alt ? g.col_scale_alt : g.col_scale;
  if (cs != nullptr) {
    const float sc = cs[c];
    val = __fmul_rn(val, (alt ? g.col_scale_invert_alt : g.col_scale_invert) ? __fsub_rn(1.f, sc) : sc);
  }
  float* dst = g.C + static_cast<long long>(r) * g.ldc + c;
  *dst = g.accumulate ? __fadd_rn(*dst, val) : val;
}

// =============================================================================================== block helpers
template <int THREADS>
__device__ __forceinline__ float block_sum(float v, float* scratch /* >= THREADS/32 floats */) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  float s = 0.f;
  if (warp == 0) {
    s = lane < THREADS / 32 ? scratch[lane] : 0.f;
    s = warp_sum(s);
  }
  __syncthreads();
  return s;   // valid in warp 0
}

constexpr int kRowThreads = 128;

// =============================================================================================== forward updates (tape)
struct UpdFwdParams {
  int N, D;
  float eps;
  const float* eps_dev;   // nullable: overrides eps
  int dir, dir_alt;
  const uint8_t* row_set;
  const float *x, *v, *g;
  const float *S, *T, *Q;
  const float *mask, *mask_alt;
  int invert, invert_alt;
  float* out;
  float* logdet;
  int32_t* flag;
  UpdExtra ex;
};

// One CTA per chain: the arithmetic of update_math.cuh with per-chain direction / mask set.
template <bool POSITION>
__global__ void __launch_bounds__(kRowThreads) train_update_fwd_kernel(const UpdFwdParams p) {
  pdl_wait();
  pdl_trigger();
  __shared__ float scratch[kRowThreads / 32];
  const int row = blockIdx.x;
  const bool alt = p.row_set != nullptr && p.row_set[row] != 0;
  const int dir = alt ? p.dir_alt : p.dir;
  const float* mask = alt ? p.mask_alt : p.mask;
  const int invert = alt ? p.invert_alt : p.invert;
  const float eps = p.eps_dev != nullptr ? *p.eps_dev : p.eps;
  const float heps = __fmul_rn(0.5f, eps);
  float ld = 0.f;
  bool bad = false;
  const float* nmask = alt ? p.ex.nmask_alt : p.ex.nmask;
  const int ninv = alt ? p.ex.ninv_alt : p.ex.ninv;
  for (int d = threadIdx.x; d < p.D; d += kRowThreads) {
    const size_t at = static_cast<size_t>(row) * p.D + d;
    const float S = p.S ? p.S[at] : 0.f, Tr = p.S ? p.T[at] : 0.f, Q = p.S ? p.Q[at] : 0.f;
    float val;
    if (POSITION) {
      const float m = invert ? __fsub_rn(1.f, mask[d]) : mask[d];
      val = dir > 0 ? position_elem<1>(p.x[at], p.v[at], S, Tr, Q, m, eps, ld, bad)
                    : position_elem<-1>(p.x[at], p.v[at], S, Tr, Q, m, eps, ld, bad);
    } else {
      val = dir > 0 ? momentum_elem<1>(p.v[at], p.g[at], S, Tr, Q, eps, heps, ld, bad)
                    : momentum_elem<-1>(p.v[at], p.g[at], S, Tr, Q, eps, heps, ld, bad);
    }
    p.out[at] = val;
    if (p.ex.copy[0] != nullptr) p.ex.copy[0][at] = val;
    if (p.ex.copy[1] != nullptr) p.ex.copy[1][at] = val;
    if (p.ex.masked != nullptr) p.ex.masked[at] = __fmul_rn(ninv ? __fsub_rn(1.f, nmask[d]) : nmask[d], val);
  }
  const float tot = block_sum<kRowThreads>(ld, scratch);
  if (threadIdx.x == 0 && p.logdet != nullptr) p.logdet[row] = __fadd_rn(p.logdet[row], tot);
  flag_nonfinite(p.flag, bad);
}

// out = m (*) x with the per-chain mask set (the second input of position_fn, materialised for the tape)
__global__ void __launch_bounds__(kRowThreads) mask_mul_sets_kernel(const float* __restrict__ x, const float* mask,
                                                                    int invert, const float* mask_alt, int invert_alt,
                                                                    const uint8_t* row_set, int D, float* out) {
  const int row = blockIdx.x;
  const bool alt = row_set != nullptr && row_set[row] != 0;
  const float* m = alt ? mask_alt : mask;
  const int inv = alt ? invert_alt : invert;
  for (int d = threadIdx.x; d < D; d += kRowThreads) {
    const size_t at = static_cast<size_t>(row) * D + d;
    out[at] = __fmul_rn(inv ? __fsub_rn(1.f, m[d]) : m[d], x[at]);
  }
}

// =============================================================================================== backward updates
struct UpdBwdParams {
  l2hmc_update_bwd_args a;
};

__device__ __forceinline__ void heads_bwd(const l2hmc_update_bwd_args& a, int row, int d, size_t at, float S, float Q,
                                          float dS, float dT, float dQ) {
  const float es = expf(a.coeff_s[d]);
  const float t = S / es;
  float* dst = a.dA + static_cast<size_t>(row) * 3 * a.D;
  dst[d] = dS * es * (1.f - t * t);
  dst[a.D + d] = dT;
  dst[2 * a.D + d] = dQ * expf(a.coeff_q[d]);
  a.ps[at] = dS * S;
  a.pq[at] = dQ * Q;
}

__global__ void __launch_bounds__(kRowThreads) momentum_bwd_kernel(const UpdBwdParams pp) {
  pdl_wait();
  pdl_trigger();
  __shared__ float scratch[kRowThreads / 32];
  const l2hmc_update_bwd_args& a = pp.a;
  const int row = blockIdx.x;
  const bool alt = a.row_set != nullptr && a.row_set[row] != 0;
  const int dir = alt ? a.dir_alt : a.dir;
  const float eps = a.eps_dev != nullptr ? *a.eps_dev : a.eps, heps = 0.5f * eps;
  const float gld = a.gld ? a.gld[row] : 0.f;
  float de = 0.f;
  for (int d = threadIdx.x; d < a.D; d += kRowThreads) {
    const size_t at = static_cast<size_t>(row) * a.D + d;
    const float gvp = a.gv[at], v = a.v[at], g = a.g[at];
    if (a.S == nullptr) {               // S = T = Q = 0: v' = v -+ eps/2 g
      const float sgn = dir > 0 ? -1.f : 1.f;
      a.dg[at] = sgn * heps * gvp;
      de += sgn * 0.5f * g * gvp;
      continue;
    }
    const float S = a.S[at], Tr = a.T[at], Q = a.Q[at];
    const float sc = S * (dir > 0 ? heps : -heps);
    const float se = expf(sc), te = expf(Q * eps);
    const float inner = te * g - Tr;
    float dv, ds, d_inner, dS;
    if (dir > 0) {
      dv = gvp * se;
      ds = gvp * v * se + gld;
      d_inner = -heps * gvp;
      de += ds * 0.5f * S - 0.5f * gvp * inner;
      dS = heps * ds;
    } else {
      const float vprime = se * (v + heps * inner);
      ds = gvp * vprime + gld;
      const float du = gvp * se;
      dv = du;
      d_inner = heps * du;
      de += -0.5f * ds * S + 0.5f * du * inner;
      dS = -heps * ds;
    }
    const float dq = d_inner * te * g;
    const float dQ = eps * dq;
    de += dq * Q;
    a.gv[at] = dv;
    a.dg[at] = d_inner * te;
    heads_bwd(a, row, d, at, S, Q, dS, -d_inner, dQ);
  }
  const float tot = block_sum<kRowThreads>(de, scratch);
  if (threadIdx.x == 0) a.deps[row] += tot;
}

__global__ void __launch_bounds__(kRowThreads) position_bwd_kernel(const UpdBwdParams pp) {
  pdl_wait();
  pdl_trigger();
  __shared__ float scratch[kRowThreads / 32];
  const l2hmc_update_bwd_args& a = pp.a;
  const int row = blockIdx.x;
  const bool alt = a.row_set != nullptr && a.row_set[row] != 0;
  const int dir = alt ? a.dir_alt : a.dir;
  const float* mask = alt ? a.mask_alt : a.mask;
  const int invert = alt ? a.invert_alt : a.invert;
  const float eps = a.eps_dev != nullptr ? *a.eps_dev : a.eps;
  const float gld = a.gld ? a.gld[row] : 0.f;
  float de = 0.f;
  for (int d = threadIdx.x; d < a.D; d += kRowThreads) {
    const size_t at = static_cast<size_t>(row) * a.D + d;
    const float gxp = a.gx[at], x = a.x[at], v = a.v[at];
    const float m = invert ? 1.f - mask[d] : mask[d];
    const float mi = 1.f - m;
    if (a.S == nullptr) {               // x' = x +- mbar eps v
      const float sgn = dir > 0 ? 1.f : -1.f;
      a.gv[at] += sgn * mi * eps * gxp;
      de += sgn * mi * v * gxp;
      continue;
    }
    const float S = a.S[at], Tr = a.T[at], Q = a.Q[at];
    const float sc = S * (dir > 0 ? eps : -eps);
    const float se = expf(sc), te = expf(Q * eps);
    const float inner = te * v + Tr;
    const float dy = mi * gxp;
    const float dx = m * gxp + dy * se;
    float ds, d_inner, dS;
    if (dir > 0) {
      ds = dy * x * se + gld * mi;
      dS = eps * ds;
      d_inner = eps * dy;
      de += ds * S + dy * inner;
    } else {
      const float y = se * (x - eps * inner);
      ds = dy * y + gld * mi;
      dS = -eps * ds;
      const float dw = dy * se;
      d_inner = -eps * dw;
      de += -ds * S - dw * inner;
    }
    const float dq = d_inner * te * v;
    const float dQ = eps * dq;
    de += dq * Q;
    a.gx[at] = dx;
    a.gv[at] += d_inner * te;
    heads_bwd(a, row, d, at, S, Q, dS, d_inner, dQ);
  }
  const float tot = block_sum<kRowThreads>(de, scratch);
  if (threadIdx.x == 0) a.deps[row] += tot;
}

// =============================================================================================== stencil adjoints
// One warp per chain, the chain staged in shared memory (any T x X).
//   HVP: out += d/dx [ g(x) . w ],  g = beta dS/dx:   c(site) = beta cos P(site) * Pw(site),
//        Pw = the plaquette stencil applied to w;  out0(t,x) += c(t,x) - c(t,x-1);  out1(t,x) += -c(t,x) + c(t-1,x)
struct HvpParams {
  const float* x;
  const float* w;
  float beta;
  int N, T, X;
  float* out;
};

__device__ __forceinline__ float plaq_at(const float* s, int i, int T, int X) {
  const int t = i / X, xx = i - t * X;
  const int sr = t * X + (xx + 1 == X ? 0 : xx + 1);
  const int sd = (t + 1 == T ? 0 : t + 1) * X + xx;
  return __fadd_rn(__fsub_rn(__fsub_rn(s[2 * i], s[2 * i + 1]), s[2 * sr]), s[2 * sd + 1]);
}

// out[link] (+)= stencil adjoint of the per-site coefficients c: (c(t,x) - c(t,x-1), -c(t,x) + c(t-1,x))
__device__ __forceinline__ void site_adjoint(const float* c, int i, int T, int X, float& o0, float& o1) {
  const int t = i / X, xx = i - t * X;
  const int sl = t * X + (xx == 0 ? X - 1 : xx - 1);
  const int su = (t == 0 ? T - 1 : t - 1) * X + xx;
  o0 = c[i] - c[sl];
  o1 = -c[i] + c[su];
}

__global__ void __launch_bounds__(128) force_hvp_kernel(const HvpParams p, int warps_per_block) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ __align__(16) float s_all[];
  const int T = p.T, X = p.X, SITES = T * X, D = 2 * SITES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* sx = s_all + static_cast<size_t>(warp) * (2 * D + SITES);
  float* sw = sx + D;
  float* sc = sw + D;
  for (int n = blockIdx.x * warps_per_block + warp; n < p.N; n += gridDim.x * warps_per_block) {
    const size_t base = static_cast<size_t>(n) * D;
    for (int i = lane; i < D; i += 32) sx[i] = p.x[base + i], sw[i] = p.w[base + i];
    __syncwarp();
    for (int i = lane; i < SITES; i += 32) {
      const float P = plaq_at(sx, i, T, X);
      const float Pw = plaq_at(sw, i, T, X);
      sc[i] = p.beta * cosf(P) * Pw;
    }
    __syncwarp();
    for (int i = lane; i < SITES; i += 32) {
      float o0, o1;
      site_adjoint(sc, i, T, X, o0, o1);
      p.out[base + 2 * i] += o0;
      p.out[base + 2 * i + 1] += o1;
    }
    __syncwarp();
  }
}

// a[n] = d loss / d (h_old - h_new + sumlogdet) through p = exp(min(., 0)), non-finite -> 0
__global__ void accept_bwd_kernel(const float* __restrict__ h_old, const float* __restrict__ h_new,
                                  const float* __restrict__ sld, const float* __restrict__ gp, int N, float* a) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float arg = __fsub_rn(h_old[n], h_new[n]);
  if (sld) arg = __fadd_rn(arg, sld[n]);
  const float pr = expf(fminf(arg, 0.f));
  a[n] = (arg <= 0.f && isfinite(pr)) ? gp[n] * pr : 0.f;   // tf.minimum routes the tie to its first argument
}

// gx = gx_post - a * force(x_M),  gv = -a * v_M      (H_new = beta S(x_M) + v_M^2 / 2 enters with a minus sign)
__global__ void __launch_bounds__(kRowThreads) traj_bwd_init_kernel(const float* __restrict__ a,
                                                                   const float* __restrict__ force,
                                                                   const float* __restrict__ v,
                                                                   const float* __restrict__ gx_post, int D, float* gx,
                                                                   float* gv) {
  const int row = blockIdx.x;
  const float ar = a[row];
  for (int d = threadIdx.x; d < D; d += kRowThreads) {
    const size_t at = static_cast<size_t>(row) * D + d;
    gx[at] = (gx_post ? gx_post[at] : 0.f) - ar * force[at];
    gv[at] = -ar * v[at];
  }
}

// =============================================================================================== loss backward
struct LossBwdParams {
  const float *x, *x_prop, *z, *px, *pz;
  int N, T, X, metric;
  float loss_scale, aux_weight, std_weight, charge_weight;
  float *g_xprop, *g_px, *g_pz;
};

__device__ __forceinline__ float metric_fwd(int metric, float a, float b) {
  switch (metric) {
    case L2HMC_METRIC_L1: return fabsf(a - b);
    case L2HMC_METRIC_L2: return (a - b) * (a - b);
    case L2HMC_METRIC_COS: return fabsf(cosf(a) - cosf(b));
    case L2HMC_METRIC_COS2: { const float d = cosf(a) - cosf(b); return d * d; }
    default: return 1.f - cosf(a - b);
  }
}
__device__ __forceinline__ float sgnf(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }
// d metric(a, b) / d b
__device__ __forceinline__ float metric_db(int metric, float a, float b) {
  switch (metric) {
    case L2HMC_METRIC_L1: return -sgnf(a - b);
    case L2HMC_METRIC_L2: return -2.f * (a - b);
    case L2HMC_METRIC_COS: return sgnf(cosf(a) - cosf(b)) * sinf(b);
    case L2HMC_METRIC_COS2: return 2.f * (cosf(a) - cosf(b)) * sinf(b);
    default: return -sinf(a - b);
  }
}
__device__ __forceinline__ float proj_approx(float p) {   // gauge_model.py:94-108
  return 2.f * sinf(p) - sinf(2.f * p) + 0.6666666666666666f * sinf(3.f * p) - 0.5f * sinf(4.f * p);
}
__device__ __forceinline__ float proj_approx_d(float p) {
  return 2.f * (cosf(p) - cosf(2.f * p) + cosf(3.f * p) - cosf(4.f * p));
}
__device__ __forceinline__ float charge_fft(const float* s, int T, int X, int lane) {
  float q = 0.f;
  for (int i = lane; i < T * X; i += 32) q += proj_approx(plaq_at(s, i, T, X));
  return warp_sum(q) / kTwoPi;
}

__global__ void __launch_bounds__(128) loss_bwd_kernel(const LossBwdParams p, int warps_per_block) {
  extern __shared__ __align__(16) float s_all[];
  const int T = p.T, X = p.X, SITES = T * X, D = 2 * SITES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* s0 = s_all + static_cast<size_t>(warp) * 3 * D;
  float* s1 = s0 + D;
  float* s2 = s1 + D;
  const float inv_n = 1.f / static_cast<float>(p.N);
  for (int n = blockIdx.x * warps_per_block + warp; n < p.N; n += gridDim.x * warps_per_block) {
    const size_t base = static_cast<size_t>(n) * D;
    float xs = 0.f, zs = 0.f;
    for (int i = lane; i < D; i += 32) {
      const float a = p.x[base + i], b = p.x_prop[base + i], c = p.z[base + i];
      s0[i] = a, s1[i] = b, s2[i] = c;
      xs += metric_fwd(p.metric, a, b);
      zs += metric_fwd(p.metric, c, b);
    }
    __syncwarp();
    xs = warp_sum(xs);
    zs = warp_sum(zs);
    const float qx = charge_fft(s0, T, X, lane), qp = charge_fft(s1, T, X, lane), qz = charge_fft(s2, T, X, lane);
    const float e = 1e-3f;
    const float px = p.px[n], pz = p.pz[n];
    const float ls = p.loss_scale, aw = p.aux_weight, sw = p.std_weight, cw = p.charge_weight;
    const float x_std = xs * px + e, z_std = aw * (zs * pz + e);
    const float dstd_dx = sw * (-ls / (x_std * x_std) - 1.f / ls);
    const float dstd_dz = sw * (-ls / (z_std * z_std) - 1.f / ls);
    if (lane == 0) {
      p.g_px[n] = inv_n * (dstd_dx * xs + cw * fabsf(qx - qp));
      p.g_pz[n] = inv_n * (dstd_dz * aw * zs + cw * aw * fabsf(qz - qp));
    }
    const float kx = inv_n * dstd_dx * px;
    const float kz = inv_n * dstd_dz * aw * pz;
    const float kq = inv_n * cw * (-px * sgnf(qx - qp) - aw * pz * sgnf(qz - qp)) / kTwoPi;
    // metric part (each lane touches only its own entries of s0 / s2)
    for (int i = lane; i < D; i += 32) {
      const float b = s1[i];
      s0[i] = kx * metric_db(p.metric, s0[i], b) + kz * metric_db(p.metric, s2[i], b);
    }
    __syncwarp();
    // charge part: per-site coefficient into s2, then the link stencil
    for (int i = lane; i < SITES; i += 32) s2[i] = kq * proj_approx_d(plaq_at(s1, i, T, X));
    __syncwarp();
    for (int i = lane; i < SITES; i += 32) {
      float o0, o1;
      site_adjoint(s2, i, T, X, o0, o1);
      p.g_xprop[base + 2 * i] = s0[2 * i] + o0;
      p.g_xprop[base + 2 * i + 1] = s0[2 * i + 1] + o1;
    }
    __syncwarp();
  }
}

template <typename K>
int warp_per_chain_launch(K kernel, size_t per_warp_bytes, int N, const char* what, int* grid, int* warps, size_t* smem) {
  int dev = 0, max_smem = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int w = 4;
  while (w > 1 && per_warp_bytes * w > 100 * 1024) w >>= 1;
  const size_t bytes = per_warp_bytes * w;
  if (bytes > static_cast<size_t>(max_smem)) {
    set_error("%s: a chain (%zu B staged) does not fit shared memory (%d B)", what, bytes, max_smem);
    return L2HMC_ERR_UNSUPPORTED;
  }
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(%s): %s", what, cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  *grid = max(1, min((N + w - 1) / w, sms * 8));
  *warps = w;
  *smem = bytes;
  return L2HMC_OK;
}

// =============================================================================================== reductions / Adam
// stage 1: partial[b] = sum over a grid-stride slice (fixed assignment); stage 2: one block, fixed tree
template <bool SQUARE>
__global__ void __launch_bounds__(256) reduce_stage1_kernel(const float* __restrict__ v, long long n, float* partial) {
  __shared__ float scratch[8];
  float s = 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256) {
    const float t = v[i];
    s += SQUARE ? t * t : t;
  }
  const float tot = block_sum<256>(s, scratch);
  if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(256) reduce_stage2_kernel(const float* __restrict__ partial, int nb, float* out) {
  __shared__ float scratch[8];
  float s = 0.f;
  for (int i = threadIdx.x; i < nb; i += 256) s += partial[i];
  const float tot = block_sum<256>(s, scratch);
  if (threadIdx.x == 0) out[0] = tot;
}

// Column sums of a row-major matrix (bias and coeff_* gradients: the batch sum of an activation gradient):
// stage 1: partial[slab][c] over a slab of rows, threads along the contiguous columns; stage 2: slabs in fixed order.
__global__ void __launch_bounds__(128) colsum_stage1_kernel(const float* __restrict__ B, long long ld, int rows, int cols,
                                                            int slabs, float* partial) {
  const int c = blockIdx.x * 128 + threadIdx.x;
  if (c >= cols) return;
  const int per = (rows + slabs - 1) / slabs;
  const int r0 = blockIdx.y * per, r1 = min(rows, r0 + per);
  float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  int r = r0;
  for (; r + 7 < r1; r += 8) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = B[static_cast<long long>(r + j) * ld + c];      // 8 independent loads in flight
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += v[j];
  }
  for (; r < r1; ++r) s[0] += B[static_cast<long long>(r) * ld + c];
  partial[static_cast<size_t>(blockIdx.y) * cols + c] = ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
}
// stage 2: a CTA owns 32 columns; warp g adds slabs g, g + 8, ... and the 8 warp sums meet in warp order
__global__ void __launch_bounds__(256) colsum_stage2_kernel(const float* __restrict__ partial, int slabs, int cols,
                                                            float* out) {
  __shared__ float sm[8][32];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < cols)
    for (int q = g; q < slabs; q += 8) s += partial[static_cast<size_t>(q) * cols + c];
  sm[g][lane] = s;
  __syncthreads();
  if (g == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += sm[j][lane];
    out[c] = t;
  }
}

// The same two stages for up to kMaxColSumJobs matrices of equal row count in one launch each (blockIdx.z / .y = job)
struct ColSumJobs {
  ColSumJob j[kMaxColSumJobs];
  int rows, slabs;
  float* partial;
  size_t pstride;      // floats of partial per job
};
__global__ void __launch_bounds__(128) colsum_multi_stage1_kernel(const ColSumJobs q) {
  const ColSumJob& jb = q.j[blockIdx.z];
  const int c = blockIdx.x * 128 + threadIdx.x;
  if (c >= jb.cols) return;
  const float* __restrict__ B = jb.B;
  const long long ld = jb.ld;
  const int per = (q.rows + q.slabs - 1) / q.slabs;
  const int r0 = blockIdx.y * per, r1 = min(q.rows, r0 + per);
  float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  int r = r0;
  for (; r + 7 < r1; r += 8) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = B[static_cast<long long>(r + j) * ld + c];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += v[j];
  }
  for (; r < r1; ++r) s[0] += B[static_cast<long long>(r) * ld + c];
  q.partial[blockIdx.z * q.pstride + static_cast<size_t>(blockIdx.y) * jb.cols + c] =
      ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
}
__global__ void __launch_bounds__(256) colsum_multi_stage2_kernel(const ColSumJobs q) {
  __shared__ float sm[8][32];
  const ColSumJob& jb = q.j[blockIdx.y];
  const float* __restrict__ partial = q.partial + blockIdx.y * q.pstride;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < jb.cols)
    for (int k = g; k < q.slabs; k += 8) s += partial[static_cast<size_t>(k) * jb.cols + c];
  sm[g][lane] = s;
  __syncthreads();
  if (g == 0 && c < jb.cols) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += sm[j][lane];
    jb.out[0][c] = t;
    if (jb.out[1]) jb.out[1][c] = t;
    if (jb.out[2]) jb.out[2][c] = t;
  }
}

// tf.train.AdamOptimizer (lr_t = lr sqrt(1 - b2^t) / (1 - b1^t) folded on the host):
//   m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;  theta -= lr_t m / (sqrt(v) + eps)
// with g scaled by clip / max(||g||, clip) when a global-norm clip is given (tf.clip_by_global_norm).
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ param, const float* __restrict__ grad,
                                                   float* __restrict__ m, float* __restrict__ v, long long n, float lr_t,
                                                   float b1, float b2, float eps, float grad_scale,
                                                   const float* sumsq, float clip) {
  float scale = grad_scale;
  if (sumsq != nullptr && clip > 0.f) {
    const float norm = sqrtf(sumsq[0]) * fabsf(grad_scale);
    scale *= clip / fmaxf(norm, clip);
  }
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256) {
    const float g = grad[i] * scale;
    const float mi = b1 * m[i] + (1.f - b1) * g;
    const float vi = b2 * v[i] + (1.f - b2) * g * g;
    m[i] = mi, v[i] = vi;
    param[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

// tf.clip_by_global_norm (gauge_model.py:826-827): every gradient scaled by clip / max(||g||, clip), in place -- the
// same product adam_kernel forms when it is handed the clip, so clipping first and stepping with no clip gives the
// same bits.
__global__ void __launch_bounds__(256) clip_kernel(float* __restrict__ grad, long long n, const float* sumsq, float clip) {
  const float scale = 1.f * (clip / fmaxf(sqrtf(sumsq[0]) * 1.f, clip));
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256)
    grad[i] = grad[i] * scale;
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

// =============================================================================================== C ABI
// Skinny products (M <= 4 output rows over a long K: the t_layer kernel's gradient, [2, tape rows] x [tape rows, H]):
// a weighted column sum -- threads along the contiguous columns of B, K in slabs, M accumulators per thread; the
// slab partials [slab][M][N] go through the split-K reduce stage like any other split product.
constexpr int kSkinnyMaxM = 4, kSkinnyMinK = 2048;
template <int MM>
__global__ void __launch_bounds__(128) skinny_gemm_kernel(const l2hmc_gemm_args g, int slabs, float* partial) {
  const int n = blockIdx.x * 128 + threadIdx.x;
  if (n >= g.N) return;
  const l2hmc_gemm_seg sg = g.seg[0];
  const int per = (sg.K + slabs - 1) / slabs;
  const int k0 = blockIdx.y * per, k1 = min(sg.K, k0 + per);
  float acc[MM];
#pragma unroll
  for (int m = 0; m < MM; ++m) acc[m] = 0.f;
  int k = k0;
  for (; k + 7 < k1; k += 8) {
    float b[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] = sg.B[static_cast<long long>(k + j) * sg.b_rs + n];      // 8 loads in flight
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int m = 0; m < MM; ++m)
        if (m < g.M) acc[m] = fmaf(__ldg(sg.A + m * sg.a_rs + static_cast<long long>(k + j) * sg.a_cs), b[j], acc[m]);
  }
  for (; k < k1; ++k) {
    const float b = sg.B[static_cast<long long>(k) * sg.b_rs + n];
#pragma unroll
    for (int m = 0; m < MM; ++m)
      if (m < g.M) acc[m] = fmaf(__ldg(sg.A + m * sg.a_rs + static_cast<long long>(k) * sg.a_cs), b, acc[m]);
  }
#pragma unroll
  for (int m = 0; m < MM; ++m)
    if (m < g.M) partial[(static_cast<size_t>(blockIdx.y) * g.M + m) * g.N + n] = acc[m];
}
static int skinny_slabs(const l2hmc_gemm_args* g) {      // 0: not a skinny product
  if (g->M > kSkinnyMaxM || g->nseg != 1 || g->seg[0].b_cs != 1 || g->seg[0].K < kSkinnyMinK) return 0;
  return std::min(128, g->seg[0].K / 32);
}

// ksplit = 0: enough K slices to put ~1.5 CTAs on every SM when the output has few tiles (the per-call GEMMs of a
// config-2 training step are 32-tile launches), each slice keeping at least two K slabs
static int auto_ksplit(const l2hmc_gemm_args* g) {
  const int tiles = ((g->M + GBM - 1) / GBM) * ((g->N + GBN - 1) / GBN);
  if (tiles >= 120) return 1;
  int kmin = g->seg[0].K;
  for (int s = 1; s < g->nseg; ++s) kmin = std::min(kmin, g->seg[s].K);
  return std::max(1, std::min(8, std::min(240 / tiles, kmin / (2 * GBK))));
}

extern "C" size_t l2hmc_sgemm_workspace_bytes(const l2hmc_gemm_args* g, int ksplit) {
  if (!g || g->nseg < 1 || g->nseg > 3) return 0;
  if (skinny_slabs(g)) ksplit = skinny_slabs(g);
  else if (ksplit == 0 && l2hmc::dense_grad_nt_ok(g)) return 0;
  if (ksplit == 0) ksplit = auto_ksplit(g);
  if (ksplit <= 1) return 0;
  return sizeof(float) * static_cast<size_t>(ksplit) * g->M * g->N;
}

extern "C" int l2hmc_sgemm(const l2hmc_gemm_args* g, int ksplit, void* workspace, size_t workspace_bytes, void* stream) {
  L2HMC_REQUIRE(g, "l2hmc_sgemm: null arguments");
  L2HMC_REQUIRE(g->M >= 0 && g->N >= 0 && g->nseg >= 1 && g->nseg <= 3, "l2hmc_sgemm: bad sizes M=%d N=%d nseg=%d", g->M,
                g->N, g->nseg);
  if (g->M == 0 || g->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(g->C, "l2hmc_sgemm: null C");
  for (int s = 0; s < g->nseg; ++s)
    L2HMC_REQUIRE(g->seg[s].A && g->seg[s].B && g->seg[s].K >= 0, "l2hmc_sgemm: bad segment %d", s);
  const int skinny = skinny_slabs(g);
  if (skinny) ksplit = skinny;
  // d act = d out . W^T with both operands K-contiguous: the chunked kernel of the forward layers (nets.cu)
  if (!skinny && ksplit == 0 && l2hmc::dense_grad_nt_ok(g)) {
    if (g->row_set) L2HMC_REQUIRE(!g->col_scale == !g->col_scale_alt, "l2hmc_sgemm: col_scale and col_scale_alt go together");
    return l2hmc::dense_grad_nt(g, static_cast<cudaStream_t>(stream));
  }
  if (ksplit == 0) ksplit = auto_ksplit(g);
  if (ksplit < 1) ksplit = 1;
  if (ksplit > 1)
    L2HMC_REQUIRE(workspace && workspace_bytes >= l2hmc_sgemm_workspace_bytes(g, ksplit),
                  "l2hmc_sgemm: split-K workspace too small (%zu < %zu bytes)", workspace_bytes,
                  l2hmc_sgemm_workspace_bytes(g, ksplit));
  if (g->row_set) L2HMC_REQUIRE(!g->col_scale == !g->col_scale_alt, "l2hmc_sgemm: col_scale and col_scale_alt go together");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (skinny) {
    skinny_gemm_kernel<kSkinnyMaxM><<<dim3((g->N + 127) / 128, skinny), 128, 0, st>>>(*g, skinny,
                                                                                     static_cast<float*>(workspace));
    int rc = check_launch("skinny_gemm_kernel");
    return rc ? rc : l2hmc::launch_splitk_reduce(static_cast<float*>(workspace), skinny, g, st);
  }
  const int cluster = ksplit >= 2 && ksplit <= 8;      // portable cluster size
  GemmParams p{*g, ksplit, static_cast<float*>(workspace), cluster};
  dim3 grid((g->N + GBN - 1) / GBN, (g->M + GBM - 1) / GBM, ksplit);
  const bool akc = g->seg[0].a_cs == 1, bkc = g->seg[0].b_rs == 1 && g->seg[0].b_cs != 1;
  bool vec4 = !akc && !bkc && g->M % 4 == 0 && g->N % 4 == 0;
  for (int s = 0; s < g->nseg && vec4; ++s) {
    const l2hmc_gemm_seg& sg = g->seg[s];
    vec4 = sg.a_rs == 1 && sg.b_cs == 1 && sg.a_cs % 4 == 0 && sg.b_rs % 4 == 0 && l2hmc::aligned16(sg.A) && l2hmc::aligned16(sg.B);
  }
  void (*kern)(const GemmParams) = akc && bkc ? sgemm_kernel<true, true>
                                   : akc      ? sgemm_kernel<true, false>
                                   : bkc      ? sgemm_kernel<false, true>
                                   : vec4     ? sgemm_kernel<false, false, true>
                                              : sgemm_kernel<false, false>;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid, cfg.blockDim = dim3(kGemmThreads), cfg.dynamicSmemBytes = 0, cfg.stream = st;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 1, attr.val.clusterDim.y = 1, attr.val.clusterDim.z = cluster ? ksplit : 1;
  cudaLaunchAttribute attrs[2];
  attrs[0] = attr;
  attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs, cfg.numAttrs = l2hmc::pdl_enabled() ? 2 : 1;
  if (cudaLaunchKernelEx(&cfg, kern, p) != cudaSuccess) return check_launch("sgemm_kernel", 0);
  int rc = check_launch("sgemm_kernel");
  if (rc || ksplit == 1 || cluster) return rc;
  return l2hmc::launch_splitk_reduce(p.partial, ksplit, g, st);
}

int l2hmc::launch_splitk_reduce(const float* partial, int ksplit, const l2hmc_gemm_args* g, cudaStream_t stream) {
  const size_t total = static_cast<size_t>(g->M) * g->N;
  launch_pdl(splitk_reduce_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0, stream, partial, ksplit, *g);
  return check_launch("splitk_reduce_kernel");
}

extern "C" int l2hmc_train_update_fwd(int position, const float* x, const float* v, const float* g, const float* S,
                                      const float* Tr, const float* Q, const float* mask, int invert,
                                      const float* mask_alt, int invert_alt, const uint8_t* row_set, float eps, int dir,
                                      int dir_alt, int N, int D, float* out, float* logdet, int32_t* nonfinite_flag,
                                      void* stream) {
  return l2hmc::train_update_fwd(position, x, v, g, S, Tr, Q, mask, invert, mask_alt, invert_alt, row_set, eps, nullptr,
                                 dir, dir_alt, N, D, out, logdet, nonfinite_flag, nullptr, stream);
}

// eps_dev (nullable) overrides eps: the step size read from device memory (l2hmc_train_args.eps_dev)
int l2hmc::train_update_fwd(int position, const float* x, const float* v, const float* g, const float* S, const float* Tr,
                            const float* Q, const float* mask, int invert, const float* mask_alt, int invert_alt,
                            const uint8_t* row_set, float eps, const float* eps_dev, int dir, int dir_alt, int N, int D,
                            float* out, float* logdet, int32_t* nonfinite_flag, const UpdExtra* extra, void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(N > 0 && D > 0 && v && out, "l2hmc_train_update_fwd: bad arguments");
  L2HMC_REQUIRE((dir == 1 || dir == -1) && (!row_set || dir_alt == 1 || dir_alt == -1), "l2hmc_train_update_fwd: dir");
  L2HMC_REQUIRE((S == nullptr) == (Tr == nullptr) && (S == nullptr) == (Q == nullptr),
                "l2hmc_train_update_fwd: S, T, Q go together");
  if (position) L2HMC_REQUIRE(x && mask && (!row_set || mask_alt), "l2hmc_train_update_fwd: position needs x and masks");
  else L2HMC_REQUIRE(g, "l2hmc_train_update_fwd: momentum needs g");
  UpdFwdParams p{N, D, eps, eps_dev, dir, dir_alt, row_set, x, v, g, S, Tr, Q, mask, mask_alt, invert, invert_alt, out, logdet,
                 nonfinite_flag, extra ? *extra : UpdExtra{}};
  if (extra && extra->masked) L2HMC_REQUIRE(extra->nmask && (!row_set || extra->nmask_alt), "train_update_fwd: next mask missing");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (position) launch_pdl(train_update_fwd_kernel<true>, dim3(N), dim3(kRowThreads), 0, st, p);
  else launch_pdl(train_update_fwd_kernel<false>, dim3(N), dim3(kRowThreads), 0, st, p);
  return check_launch("train_update_fwd_kernel");
}

extern "C" int l2hmc_mask_mul_sets(const float* x, const float* mask, int invert, const float* mask_alt, int invert_alt,
                                   const uint8_t* row_set, int N, int D, float* out, void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(x && mask && out && N > 0 && D > 0 && (!row_set || mask_alt), "l2hmc_mask_mul_sets: bad arguments");
  mask_mul_sets_kernel<<<N, kRowThreads, 0, static_cast<cudaStream_t>(stream)>>>(x, mask, invert, mask_alt, invert_alt,
                                                                               row_set, D, out);
  return check_launch("mask_mul_sets_kernel");
}

extern "C" int l2hmc_update_bwd(const l2hmc_update_bwd_args* a, void* stream) {
  L2HMC_REQUIRE(a, "l2hmc_update_bwd: null arguments");
  if (a->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(a->N > 0 && a->D > 0 && a->v && a->gv && a->deps, "l2hmc_update_bwd: bad arguments");
  L2HMC_REQUIRE((a->dir == 1 || a->dir == -1) && (!a->row_set || a->dir_alt == 1 || a->dir_alt == -1),
                "l2hmc_update_bwd: dir");
  if (a->S) L2HMC_REQUIRE(a->T && a->Q && a->coeff_s && a->coeff_q && a->dA && a->ps && a->pq,
                          "l2hmc_update_bwd: the network outputs need T, Q, coeff_*, dA, ps, pq");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  UpdBwdParams p{*a};
  if (a->position) {
    L2HMC_REQUIRE(a->x && a->gx && a->mask && (!a->row_set || a->mask_alt), "l2hmc_update_bwd: position needs x, gx, masks");
    launch_pdl(position_bwd_kernel, dim3(a->N), dim3(kRowThreads), 0, st, p);
    return check_launch("position_bwd_kernel");
  }
  L2HMC_REQUIRE(a->g && a->dg, "l2hmc_update_bwd: momentum needs g and dg");
  launch_pdl(momentum_bwd_kernel, dim3(a->N), dim3(kRowThreads), 0, st, p);
  return check_launch("momentum_bwd_kernel");
}

extern "C" int l2hmc_u1_force_hvp(const float* x, const float* w, float beta, int N, int T, int X, float* out,
                                  void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(x && w && out && N > 0 && T > 0 && X > 0, "l2hmc_u1_force_hvp: bad arguments");
  int grid, warps;
  size_t smem;
  int rc = warp_per_chain_launch(force_hvp_kernel, sizeof(float) * 5u * T * X, N, "force_hvp_kernel", &grid, &warps, &smem);
  if (rc) return rc;
  HvpParams p{x, w, beta, N, T, X, out};
  launch_pdl(force_hvp_kernel, dim3(grid), dim3(warps * 32), smem, static_cast<cudaStream_t>(stream), p, warps);
  return check_launch("force_hvp_kernel");
}

extern "C" int l2hmc_accept_prob_bwd(const float* h_old, const float* h_new, const float* sumlogdet, const float* g_p,
                                     int N, float* a, void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(h_old && h_new && g_p && a && N > 0, "l2hmc_accept_prob_bwd: bad arguments");
  accept_bwd_kernel<<<(N + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(h_old, h_new, sumlogdet, g_p, N, a);
  return check_launch("accept_bwd_kernel");
}

extern "C" int l2hmc_traj_bwd_init(const float* a, const float* force, const float* v, const float* gx_post, int N, int D,
                                   float* gx, float* gv, void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(a && force && v && gx && gv && N > 0 && D > 0, "l2hmc_traj_bwd_init: bad arguments");
  traj_bwd_init_kernel<<<N, kRowThreads, 0, static_cast<cudaStream_t>(stream)>>>(a, force, v, gx_post, D, gx, gv);
  return check_launch("traj_bwd_init_kernel");
}

extern "C" int l2hmc_u1_loss_bwd(const l2hmc_loss_params* prm, const float* x, const float* x_prop, const float* z,
                                 const float* px, const float* pz, int N, int T, int X, float* g_xprop, float* g_px,
                                 float* g_pz, void* stream) {
  L2HMC_REQUIRE(prm, "l2hmc_u1_loss_bwd: null parameters");
  L2HMC_REQUIRE(N > 0 && T > 0 && X > 0, "l2hmc_u1_loss_bwd: bad sizes N=%d T=%d X=%d", N, T, X);
  L2HMC_REQUIRE(x && x_prop && z && px && pz && g_xprop && g_px && g_pz, "l2hmc_u1_loss_bwd: null pointer");
  L2HMC_REQUIRE(prm->metric >= L2HMC_METRIC_L1 && prm->metric <= L2HMC_METRIC_COS_DIFF, "l2hmc_u1_loss_bwd: metric %d",
                prm->metric);
  int grid, warps;
  size_t smem;
  int rc = warp_per_chain_launch(loss_bwd_kernel, sizeof(float) * 6u * T * X, N, "loss_bwd_kernel", &grid, &warps, &smem);
  if (rc) return rc;
  LossBwdParams p{x, x_prop, z, px, pz, N, T, X, prm->metric, prm->loss_scale, prm->aux_weight, prm->std_weight,
                  prm->charge_weight, g_xprop, g_px, g_pz};
  loss_bwd_kernel<<<grid, warps * 32, smem, static_cast<cudaStream_t>(stream)>>>(p, warps);
  return check_launch("loss_bwd_kernel");
}

// enough slabs that the loads in flight cover HBM latency (40 960 tape rows: 512 slabs x 8 column tiles of CTAs,
// 8 loads per thread); the second stage splits the slabs over 8 warps
static int colsum_slabs(int rows) { return std::max(1, std::min(512, rows / 32)); }

extern "C" size_t l2hmc_col_sum_workspace_bytes(int rows, int cols) {
  if (rows <= 0 || cols <= 0) return 0;
  return sizeof(float) * static_cast<size_t>(colsum_slabs(rows)) * cols;
}

extern "C" int l2hmc_col_sum(const float* B, long long ld, int rows, int cols, float* out, void* workspace,
                             size_t workspace_bytes, void* stream) {
  L2HMC_REQUIRE(B && out && rows > 0 && cols > 0 && ld >= cols, "l2hmc_col_sum: bad arguments");
  L2HMC_REQUIRE(workspace && workspace_bytes >= l2hmc_col_sum_workspace_bytes(rows, cols),
                "l2hmc_col_sum: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int slabs = colsum_slabs(rows);
  float* partial = static_cast<float*>(workspace);
  colsum_stage1_kernel<<<dim3((cols + 127) / 128, slabs), 128, 0, st>>>(B, ld, rows, cols, slabs, partial);
  colsum_stage2_kernel<<<(cols + 31) / 32, 256, 0, st>>>(partial, slabs, cols, out);
  return check_launch("colsum_kernel", 2);
}

size_t l2hmc::col_sum_multi_workspace_bytes(int rows, int max_cols, int njobs) {
  if (rows <= 0 || max_cols <= 0 || njobs <= 0) return 0;
  return sizeof(float) * static_cast<size_t>(colsum_slabs(rows)) * max_cols * njobs;
}

int l2hmc::col_sum_multi(const ColSumJob* jobs, int njobs, int rows, void* workspace, size_t workspace_bytes,
                         cudaStream_t st) {
  L2HMC_REQUIRE(jobs && njobs >= 1 && njobs <= kMaxColSumJobs && rows > 0, "col_sum_multi: bad arguments");
  ColSumJobs q{};
  int max_cols = 0;
  for (int i = 0; i < njobs; ++i) {
    L2HMC_REQUIRE(jobs[i].B && jobs[i].out[0] && jobs[i].cols > 0 && jobs[i].ld >= jobs[i].cols, "col_sum_multi: bad job %d", i);
    q.j[i] = jobs[i];
    max_cols = std::max(max_cols, jobs[i].cols);
  }
  L2HMC_REQUIRE(workspace && workspace_bytes >= col_sum_multi_workspace_bytes(rows, max_cols, njobs),
                "col_sum_multi: workspace too small");
  q.rows = rows, q.slabs = colsum_slabs(rows);
  q.partial = static_cast<float*>(workspace);
  q.pstride = static_cast<size_t>(q.slabs) * max_cols;
  colsum_multi_stage1_kernel<<<dim3((max_cols + 127) / 128, q.slabs, njobs), 128, 0, st>>>(q);
  colsum_multi_stage2_kernel<<<dim3((max_cols + 31) / 32, njobs), 256, 0, st>>>(q);
  return check_launch("colsum_multi_kernel", 2);
}

// ----------------------------------------------------------------------------------------------- no-allocation variant
// l2hmc_col_sum for callers that cannot provide (or allocate) a workspace: the first-stage partials live in a
// file-scope device array that every GPU loading the library holds until exit.  The array's NAME is only a host-side
// handle: its device address is looked up once per device with cudaGetSymbolAddress and cached; the kernels get that
// pointer.  Nothing has to start at zero (stage 1 writes every partial[slab][c] stage 2 reads), but all calls share
// the array, so two calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) float g_colsum_static_scratch[L2HMC_COL_SUM_STATIC_MAX_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<float*> g_colsum_static_ptr[kMaxDevices];   // zero-initialised: nullptr = not looked up on that device yet

int colsum_static_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("l2hmc_col_sum_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return L2HMC_ERR_CUDA;
  }
  float* p = g_colsum_static_ptr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {   // once per device, with that device current (racing threads look up the same address)
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_colsum_static_scratch);
    if (e != cudaSuccess || sym == nullptr) {
      set_error("l2hmc_col_sum_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
    p = static_cast<float*>(sym);
    g_colsum_static_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return L2HMC_OK;
}
}  // namespace

extern "C" int l2hmc_col_sum_static(const float* B, long long ld, int rows, int cols, float* out, void* stream) {
  L2HMC_REQUIRE(B && out && rows > 0 && cols > 0 && ld >= cols, "l2hmc_col_sum_static: bad arguments");
  const int slabs = colsum_slabs(rows);
  const size_t need = static_cast<size_t>(slabs) * cols;
  L2HMC_REQUIRE(need <= static_cast<size_t>(L2HMC_COL_SUM_STATIC_MAX_FLOATS),
                "l2hmc_col_sum_static: rows=%d x cols=%d needs %zu scratch floats (%d row slabs x cols), the built-in "
                "scratch holds L2HMC_COL_SUM_STATIC_MAX_FLOATS = %d; use l2hmc_col_sum with a caller workspace",
                rows, cols, need, slabs, L2HMC_COL_SUM_STATIC_MAX_FLOATS);
  float* partial = nullptr;
  int rc = colsum_static_scratch(&partial);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  colsum_stage1_kernel<<<dim3((cols + 127) / 128, slabs), 128, 0, st>>>(B, ld, rows, cols, slabs, partial);
  colsum_stage2_kernel<<<(cols + 31) / 32, 256, 0, st>>>(partial, slabs, cols, out);
  return check_launch("colsum_kernel (static scratch)", 2);
}

extern "C" int l2hmc_reduce_sum(const float* v, long long n, int square, float* out, float* scratch /* [1024] */,
                                void* stream) {
  L2HMC_REQUIRE(v && out && scratch && n >= 0, "l2hmc_reduce_sum: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = static_cast<int>(std::max<long long>(1, std::min<long long>(1024, (n + 255) / 256)));
  if (square) reduce_stage1_kernel<true><<<nb, 256, 0, st>>>(v, n, scratch);
  else reduce_stage1_kernel<false><<<nb, 256, 0, st>>>(v, n, scratch);
  int rc = check_launch("reduce_stage1_kernel");
  if (rc) return rc;
  reduce_stage2_kernel<<<1, 256, 0, st>>>(scratch, nb, out);
  return check_launch("reduce_stage2_kernel");
}

extern "C" int l2hmc_clip_by_global_norm(float* grad, long long n, const float* grad_sumsq, float clip_norm, void* stream) {
  if (n == 0) return L2HMC_OK;
  L2HMC_REQUIRE(grad && grad_sumsq && n > 0 && clip_norm > 0.f, "l2hmc_clip_by_global_norm: bad arguments");
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int nb = static_cast<int>(std::max<long long>(1, std::min<long long>(sms * 8LL, (n + 255) / 256)));
  clip_kernel<<<nb, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad, n, grad_sumsq, clip_norm);
  return check_launch("clip_kernel");
}

extern "C" int l2hmc_adam_step(float* param, const float* grad, float* m, float* v, long long n, float lr_t, float beta1,
                               float beta2, float epsilon, float grad_scale, const float* grad_sumsq, float clip_norm,
                               void* stream) {
  if (n == 0) return L2HMC_OK;
  L2HMC_REQUIRE(param && grad && m && v && n > 0, "l2hmc_adam_step: bad arguments");
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int nb = static_cast<int>(std::max<long long>(1, std::min<long long>(sms * 8LL, (n + 255) / 256)));
  adam_kernel<<<nb, 256, 0, static_cast<cudaStream_t>(stream)>>>(param, grad, m, v, n, lr_t, beta1, beta2, epsilon,
                                                               grad_scale, grad_sumsq, clip_norm);
  return check_launch("adam_kernel");
}
