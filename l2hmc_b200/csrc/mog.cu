// MoG / generic-target L2HMC sampler (BASELINE config 5): one thread per chain, a whole MC step per
// launch, both S/T/F networks resident in shared memory.
//
// Replaces (reference l2hmc/...):
//   utils/distributions.py:32-39,124-158  GMM energy (row-wise instead of the N x N diag_part form)
//   utils/dynamics.py:120-319             _forward_step / _backward_step / forward / backward / p_accept
//   utils/network.py:89-114,359-454       network(): embed_1..3 -> relu -> linear_1 -> relu -> S,T,F heads
//   utils/sampler.py:28-59                propose + tf_accept (p - u >= 0)
//   utils/trajectories.py:39-60           nearest-mean assignment (for the tunneling rate)
//
// Why not GEMMs: the state is k = 2 floats per chain and the nets are 2 -> 50 -> 50 -> 2; with 10^6
// independent chains the natural decomposition is a chain per thread with the weights broadcast
// from shared memory (every lane reads the same weight: one LDS.128 wavefront per four FMAs) and the
// hidden vector in registers.  HBM traffic is 16 B per chain per MC step; the kernel is FFMA bound.
// Dimensions are padded to compile-time sizes (k -> KD in {2,4,8}, H -> HP in {16,32,52,64}) with zero
// weights, which leaves every result unchanged (padded S = tanh(0) = 0, padded v = 0).
#include <algorithm>

#include "common.cuh"
#include "update_math.cuh"

namespace l2hmc {
namespace {

constexpr int kMogThreads = 128;
constexpr int kMaxMix = 16, kMaxL = 32;

struct MogParams {
  l2hmc_mog_args a;
  int k, m, H;
  const float* mus;
  const float* i_sigmas;
  const float* log_consts;
  // raw weight pointers of both nets (index 0 = XNet, 1 = VNet)
  l2hmc_net_weights net[2];
};

// shared-memory image of one network, padded: W1[HP/2][2*KD][2] (in0 | in1 columns, unit pairs interleaved), b1e[L][HP] (bias with the
// time embedding folded in, one row per step), W2t[HP][HP] (row j = weights INTO unit j), b2[HP],
// W3[HP][3*KD] (S | T | F columns per hidden unit), b3[3*KD], escale[2*KD] = exp(coeff_s | coeff_f)
template <int KD, int HP>
struct NetSmem {
  static constexpr int W1 = 0;
  static constexpr int B1E = W1 + HP * 2 * KD;
  static constexpr int W2 = B1E + kMaxL * HP;
  static constexpr int B2 = W2 + HP * HP;
  static constexpr int W3 = B2 + HP;
  static constexpr int B3 = W3 + HP * 3 * KD;
  static constexpr int ESC = B3 + 3 * KD;
  static constexpr int SIZE = (ESC + 2 * KD + 3) & ~3;
};

template <int KD, int HP>
__device__ void load_net(float* s, const l2hmc_net_weights& w, int k, int H, int L) {
  using NS = NetSmem<KD, HP>;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < NS::SIZE; i += nt) s[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < H * k; i += nt) {
    const int j = i / k, c = i % k;                       // hidden unit j, input component c
    // pairs of hidden units (2q, 2q+1) interleaved per input component: [q][c'][parity]
    s[NS::W1 + ((j >> 1) * 2 * KD + c) * 2 + (j & 1)] = w.w_in0[c * H + j];      // Keras (in, out)
    s[NS::W1 + ((j >> 1) * 2 * KD + KD + c) * 2 + (j & 1)] = w.w_in1[c * H + j];
    s[NS::W3 + j * 3 * KD + c] = w.w_s[j * k + c];
    s[NS::W3 + j * 3 * KD + KD + c] = w.w_tr[j * k + c];
    s[NS::W3 + j * 3 * KD + 2 * KD + c] = w.w_q[j * k + c];
  }
  for (int i = tid; i < H * H; i += nt) {
    const int j = i / H, kk = i % H;
    s[NS::W2 + j * HP + kk] = w.w_h[kk * H + j];
  }
  for (int i = tid; i < H; i += nt) s[NS::B2 + i] = w.b_h[i];
  for (int i = tid; i < L * H; i += nt) {
    const int step = i / H, j = i % H;
    // [cos, sin](2 pi step / L) in fp32 like the graph (dynamics.py:106-111)
    const float arg = __fdiv_rn(__fmul_rn(kTwoPi, static_cast<float>(step)), static_cast<float>(L));
    const float tt = __fadd_rn(__fadd_rn(__fmul_rn(cosf(arg), w.w_t[j]), __fmul_rn(sinf(arg), w.w_t[H + j])), w.b_t[j]);
    s[NS::B1E + step * HP + j] = __fadd_rn(__fadd_rn(w.b_in0[j], w.b_in1[j]), tt);
  }
  for (int i = tid; i < k; i += nt) {
    s[NS::B3 + i] = w.b_s[i];
    s[NS::B3 + KD + i] = w.b_tr[i];
    s[NS::B3 + 2 * KD + i] = w.b_q[i];
    s[NS::ESC + i] = expf(w.coeff_s[i]);
    s[NS::ESC + KD + i] = expf(w.coeff_q[i]);
  }
  __syncthreads();
}

// S, T, F = net([in0, in1, t(step)])   (network.py:89-114)
// Packed fp32x2 arithmetic (FFMA2: two FMAs per lane per instruction -- the only way to the full FP32 rate
// on sm_100): layer 1 pairs two hidden units, layer 2 pairs two consecutive inputs of one unit (even / odd
// partial sums, added at the end), the heads pair two outputs.  Each lane of a pair is an ordinary fmaf.
template <int KD, int HP>
__device__ __forceinline__ void net_eval(const float* __restrict__ s, int step, const float* in0, const float* in1,
                                         float* S, float* Tr, float* F) {
  using NS = NetSmem<KD, HP>;
  f32x2 in0p[KD], in1p[KD];
#pragma unroll
  for (int c = 0; c < KD; ++c) in0p[c] = pack2(in0[c], in0[c]), in1p[c] = pack2(in1[c], in1[c]);
  f32x2 h1[HP / 2];                       // (h1[2q], h1[2q+1])
#pragma unroll
  for (int q = 0; q < HP / 2; ++q) {
    f32x2 acc = *reinterpret_cast<const f32x2*>(s + NS::B1E + step * HP + 2 * q);
    const f32x2* w = reinterpret_cast<const f32x2*>(s + NS::W1 + q * 4 * KD);   // [c' = 0 .. 2KD) pairs over (2q, 2q+1)
#pragma unroll
    for (int c = 0; c < KD; ++c) {
      acc = fma2(in0p[c], w[c], acc);
      acc = fma2(in1p[c], w[KD + c], acc);
    }
    h1[q] = pack2(fmaxf(lo2(acc), 0.f), fmaxf(hi2(acc), 0.f));
  }
  f32x2 o[3 * KD / 2];
#pragma unroll
  for (int c = 0; c < 3 * KD / 2; ++c) o[c] = *reinterpret_cast<const f32x2*>(s + NS::B3 + 2 * c);
#pragma unroll 2
  for (int j = 0; j < HP; ++j) {
    f32x2 acc = pack2(s[NS::B2 + j], 0.f);
    const ulonglong2* wrow = reinterpret_cast<const ulonglong2*>(s + NS::W2 + j * HP);
#pragma unroll
    for (int q = 0; q < HP / 4; ++q) {
      const ulonglong2 w4 = wrow[q];
      acc = fma2(h1[2 * q], w4.x, acc);
      acc = fma2(h1[2 * q + 1], w4.y, acc);
    }
    const float h2 = fmaxf(__fadd_rn(lo2(acc), hi2(acc)), 0.f);
    const f32x2 h2p = pack2(h2, h2);
    const f32x2* w3 = reinterpret_cast<const f32x2*>(s + NS::W3 + j * 3 * KD);
#pragma unroll
    for (int c = 0; c < 3 * KD / 2; ++c) o[c] = fma2(h2p, w3[c], o[c]);
  }
  float of[3 * KD];
#pragma unroll
  for (int c = 0; c < 3 * KD / 2; ++c) of[2 * c] = lo2(o[c]), of[2 * c + 1] = hi2(o[c]);
#pragma unroll
  for (int c = 0; c < KD; ++c) {
    S[c] = __fmul_rn(s[NS::ESC + c], tanhf(of[c]));
    Tr[c] = of[KD + c];
    F[c] = __fmul_rn(s[NS::ESC + KD + c], tanhf(of[2 * KD + c]));
  }
}

// GMM parameters in shared memory: mus [m][KD], S [m][KD][KD] (symmetrised), logc [m]
template <int KD>
struct GmmSmem {
  static constexpr int MU = 0;
  static constexpr int SG = MU + kMaxMix * KD;
  static constexpr int LC = SG + kMaxMix * KD * KD;
  static constexpr int SIZE = (LC + kMaxMix + 3) & ~3;
};

template <int KD>
__device__ void load_gmm(float* s, const MogParams& p) {
  using GS = GmmSmem<KD>;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < GS::SIZE; i += nt) s[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < p.m * p.k; i += nt) s[GS::MU + (i / p.k) * KD + i % p.k] = p.mus[i];
  for (int i = tid; i < p.m * p.k * p.k; i += nt) {
    const int mi = i / (p.k * p.k), r = (i / p.k) % p.k, c = i % p.k;
    // d(0.5 d^T S d)/dd = 0.5 (S + S^T) d: keep the symmetric part (S is symmetric up to rounding)
    s[GS::SG + (mi * KD + r) * KD + c] =
        __fmul_rn(0.5f, __fadd_rn(p.i_sigmas[(mi * p.k + r) * p.k + c], p.i_sigmas[(mi * p.k + c) * p.k + r]));
  }
  for (int i = tid; i < p.m; i += nt) s[GS::LC + i] = p.log_consts[i];
  __syncthreads();
}

// E(x)/T and (optionally) its gradient.  distributions.py:151-158 + dynamics.py:227-242.
template <int KD, bool GRAD>
__device__ __forceinline__ float gmm_energy(const float* __restrict__ s, int m, float temp, const float* x,
                                            float* grad) {
  using GS = GmmSmem<KD>;
  float a[kMaxMix];
  float amax = -INFINITY;
  for (int i = 0; i < m; ++i) {
    float d[KD], quad = 0.f;
#pragma unroll
    for (int c = 0; c < KD; ++c) d[c] = __fsub_rn(x[c], s[GS::MU + i * KD + c]);
#pragma unroll
    for (int r = 0; r < KD; ++r) {
      float sd = 0.f;
#pragma unroll
      for (int c = 0; c < KD; ++c) sd = fmaf(s[GS::SG + (i * KD + r) * KD + c], d[c], sd);
      quad = fmaf(d[r], sd, quad);
    }
    a[i] = __fadd_rn(-__fmul_rn(0.5f, quad), s[GS::LC + i]);
    amax = fmaxf(amax, a[i]);
  }
  float z = 0.f;
  for (int i = 0; i < m; ++i) z += expf(__fsub_rn(a[i], amax));
  if (GRAD) {
#pragma unroll
    for (int c = 0; c < KD; ++c) grad[c] = 0.f;
    for (int i = 0; i < m; ++i) {
      const float wgt = __fdiv_rn(__fdiv_rn(expf(__fsub_rn(a[i], amax)), z), temp);   // softmax_i / T
#pragma unroll
      for (int r = 0; r < KD; ++r) {
        float sd = 0.f;
#pragma unroll
        for (int c = 0; c < KD; ++c)
          sd = fmaf(s[GS::SG + (i * KD + r) * KD + c], __fsub_rn(x[c], s[GS::MU + i * KD + c]), sd);
        grad[r] = fmaf(wgt, sd, grad[r]);
      }
    }
  }
  return __fdiv_rn(-__fadd_rn(amax, logf(z)), temp);
}

template <int KD, int HP>
__global__ void __launch_bounds__(kMogThreads) mog_mc_kernel(const MogParams p) {
  extern __shared__ __align__(16) float sm[];
  using NS = NetSmem<KD, HP>;
  using GS = GmmSmem<KD>;
  float* s_gmm = sm;
  float* s_x = sm + GS::SIZE;
  float* s_v = s_x + NS::SIZE;
  float* s_mask = s_v + NS::SIZE;             // [L][KD]
  const l2hmc_mog_args& a = p.a;
  const int k = p.k, L = a.L;
  load_gmm<KD>(s_gmm, p);
  if (!a.hmc) {
    load_net<KD, HP>(s_x, p.net[0], k, p.H, L);
    load_net<KD, HP>(s_v, p.net[1], k, p.H, L);
  }
  for (int i = threadIdx.x; i < L * KD; i += blockDim.x) {
    const int c = i % KD;
    s_mask[i] = c < k ? a.mask[(i / KD) * k + c] : 1.f;
  }
  __syncthreads();

  const float eps = a.eps, heps = __fmul_rn(0.5f, a.eps), temp = a.temperature;
  const bool rng = a.v_f == nullptr;
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += gridDim.x * blockDim.x) {
    const unsigned long long chain = a.chain0 + static_cast<unsigned long long>(n);
    float ud = 1.f, ua = 0.f;
    if (rng) philox_uniform2(a.seed, chain, a.mc_step, ud, ua);
    bool fwd = true;
    if (!a.hmc) {
      if (a.dir_mask) fwd = a.dir_mask[n] > 0.5f;
      else if (rng) fwd = ud >= 0.5f;
    }
    bool have_acc = false;
    float u_acc = 0.f;
    if (a.u_acc) u_acc = a.u_acc[n], have_acc = true;
    else if (rng) u_acc = ua, have_acc = true;

    float x[KD], v[KD], x0[KD];
#pragma unroll
    for (int c = 0; c < KD; ++c) x[c] = c < k ? a.x[static_cast<size_t>(n) * k + c] : 0.f;
    if (rng) {
#pragma unroll
      for (int q = 0; q < (KD + 3) / 4; ++q) {
        float z[4];
        philox_normal4(a.seed, chain, a.mc_step, static_cast<uint32_t>(q), z);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (4 * q + j < KD) v[4 * q + j] = 4 * q + j < k ? z[j] : 0.f;
      }
    } else {
      const float* src = (fwd || a.v_b == nullptr) ? a.v_f : a.v_b;
#pragma unroll
      for (int c = 0; c < KD; ++c) v[c] = c < k ? src[static_cast<size_t>(n) * k + c] : 0.f;
    }
    float ke0 = 0.f;
#pragma unroll
    for (int c = 0; c < KD; ++c) x0[c] = x[c], ke0 = __fadd_rn(ke0, __fmul_rn(v[c], v[c]));
    const float h0 = __fadd_rn(gmm_energy<KD, false>(s_gmm, p.m, temp, x, nullptr), __fmul_rn(0.5f, ke0));

    float logjac = 0.f;
    bool bad = false;
    float S[KD], Tr[KD], F[KD], g[KD], in1[KD];
#pragma unroll
    for (int c = 0; c < KD; ++c) S[c] = Tr[c] = F[c] = 0.f;
    auto momentum = [&](int step) {
      gmm_energy<KD, true>(s_gmm, p.m, temp, x, g);
      if (!a.hmc) net_eval<KD, HP>(s_v, step, x, g, S, Tr, F);
#pragma unroll
      for (int c = 0; c < KD; ++c)
        v[c] = fwd ? momentum_elem<1>(v[c], g[c], S[c], Tr[c], F[c], eps, heps, logjac, bad)
                   : momentum_elem<-1>(v[c], g[c], S[c], Tr[c], F[c], eps, heps, logjac, bad);
    };
    auto position = [&](int step, bool invert) {
      float mk[KD];
#pragma unroll
      for (int c = 0; c < KD; ++c) {
        const float m = s_mask[step * KD + c];
        mk[c] = invert ? __fsub_rn(1.f, m) : m;
        in1[c] = __fmul_rn(mk[c], x[c]);
      }
      if (!a.hmc) net_eval<KD, HP>(s_x, step, v, in1, S, Tr, F);
#pragma unroll
      for (int c = 0; c < KD; ++c)
        x[c] = fwd ? position_elem<1>(x[c], v[c], S[c], Tr[c], F[c], mk[c], eps, logjac, bad)
                   : position_elem<-1>(x[c], v[c], S[c], Tr[c], F[c], mk[c], eps, logjac, bad);
    };
    for (int t = 0; t < L; ++t) {
      const int step = fwd ? t : L - 1 - t;
      momentum(step);
      position(step, !fwd);      // forward keeps m first; backward keeps 1 - m first
      position(step, fwd);
      momentum(step);
    }
    float ke1 = 0.f;
#pragma unroll
    for (int c = 0; c < KD; ++c) ke1 = __fadd_rn(ke1, __fmul_rn(v[c], v[c]));
    const float h1 = __fadd_rn(gmm_energy<KD, false>(s_gmm, p.m, temp, x, nullptr), __fmul_rn(0.5f, ke1));
    const float d = __fadd_rn(__fsub_rn(h0, h1), logjac);
    float prob = expf(fminf(d, 0.f));
    if (!isfinite(prob) || isnan(d)) prob = 0.f;       // tf.where(is_finite(p), p, 0) (dynamics.py:312-319)
    const bool acc = have_acc ? (__fsub_rn(prob, u_acc) >= 0.f) : true;
    if (a.px) a.px[n] = prob;
    if (a.accepted) a.accepted[n] = acc ? 1 : 0;
    float xo[KD];
#pragma unroll
    for (int c = 0; c < KD; ++c) xo[c] = acc ? x[c] : x0[c];
#pragma unroll
    for (int c = 0; c < KD; ++c) {
      if (c < k) {
        const size_t at = static_cast<size_t>(n) * k + c;
        if (a.Lx) a.Lx[at] = x[c];
        if (a.Lv) a.Lv[at] = v[c];
        if (a.x_out) a.x_out[at] = xo[c];
      }
    }
    if (a.mode) {
      int best = 0;
      float bd = INFINITY;
      for (int i = 0; i < p.m; ++i) {
        float dd = 0.f;
#pragma unroll
        for (int c = 0; c < KD; ++c) {
          const float e = __fsub_rn(xo[c], s_gmm[GS::MU + i * KD + c]);
          dd = fmaf(e, e, dd);
        }
        if (dd < bd) bd = dd, best = i;
      }
      a.mode[n] = best;
    }
    if (bad && a.nonfinite_flag) atomicOr(a.nonfinite_flag, 1);
  }
}

template <int KD>
__global__ void __launch_bounds__(kMogThreads) gmm_kernel(const MogParams p, const float* x, float temperature, int N,
                                                          float* energy, float* force) {
  extern __shared__ __align__(16) float sm[];
  load_gmm<KD>(sm, p);
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
    float xx[KD], g[KD];
#pragma unroll
    for (int c = 0; c < KD; ++c) xx[c] = c < p.k ? x[static_cast<size_t>(n) * p.k + c] : 0.f;
    const float e = gmm_energy<KD, true>(sm, p.m, temperature, xx, g);
    if (energy) energy[n] = e;
    if (force) {
#pragma unroll
      for (int c = 0; c < KD; ++c)
        if (c < p.k) force[static_cast<size_t>(n) * p.k + c] = g[c];
    }
  }
}

int sm_count() {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}

template <int KD, int HP>
int launch_mog(const MogParams& p, cudaStream_t st) {
  const size_t smem = sizeof(float) * (GmmSmem<KD>::SIZE + 2 * NetSmem<KD, HP>::SIZE + kMaxL * KD);
  auto kern = mog_mc_kernel<KD, HP>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(mog_mc_kernel, %zu B): %s", smem, cudaGetErrorString(e));
    return L2HMC_ERR_CUDA;
  }
  int bps = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, kMogThreads, smem);
  if (bps < 1) {
    set_error("mog_mc_kernel does not fit an SM (%zu B shared)", smem);
    return L2HMC_ERR_CUDA;
  }
  const int grid = std::max(1, std::min((p.a.N + kMogThreads - 1) / kMogThreads, bps * sm_count()));
  kern<<<grid, kMogThreads, smem, st>>>(p);
  return check_launch("mog_mc_kernel");
}

template <int KD>
int launch_mog_h(const MogParams& p, cudaStream_t st) {
  if (p.H <= 16) return launch_mog<KD, 16>(p, st);
  if (p.H <= 32) return launch_mog<KD, 32>(p, st);
  if (p.H <= 52) return launch_mog<KD, 52>(p, st);
  return launch_mog<KD, 64>(p, st);
}

int fill_gmm(const l2hmc_gmm* g, MogParams& p) {
  L2HMC_REQUIRE(g && g->mus && g->i_sigmas && g->log_consts, "l2hmc_gmm: null pointer");
  L2HMC_REQUIRE(g->k >= 1 && g->k <= 8 && g->m >= 1 && g->m <= kMaxMix, "l2hmc_gmm: need 1 <= k <= 8, 1 <= m <= %d",
                kMaxMix);
  p.k = g->k, p.m = g->m, p.mus = g->mus, p.i_sigmas = g->i_sigmas, p.log_consts = g->log_consts;
  return L2HMC_OK;
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_gmm_energy_force(const l2hmc_gmm* gmm, const float* x, float temperature, int N, float* energy,
                                      float* force, void* stream) {
  MogParams p{};
  int rc = fill_gmm(gmm, p);
  if (rc) return rc;
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(x && (energy || force) && N > 0 && temperature > 0.f, "l2hmc_gmm_energy_force: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = std::max(1, std::min((N + kMogThreads - 1) / kMogThreads, 8 * sm_count()));
#define L2HMC_GMM(KD)                                                                                      \
  gmm_kernel<KD><<<grid, kMogThreads, sizeof(float) * GmmSmem<KD>::SIZE, st>>>(p, x, temperature, N, energy, force)
  if (p.k <= 2) L2HMC_GMM(2);
  else if (p.k <= 4) L2HMC_GMM(4);
  else L2HMC_GMM(8);
#undef L2HMC_GMM
  return check_launch("gmm_kernel");
}

extern "C" int l2hmc_mog_mc_step(const l2hmc_gmm* gmm, const l2hmc_mog_args* a, void* stream) {
  MogParams p{};
  int rc = fill_gmm(gmm, p);
  if (rc) return rc;
  L2HMC_REQUIRE(a, "l2hmc_mog_mc_step: null args");
  if (a->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(a->x && a->mask && a->N > 0, "l2hmc_mog_mc_step: null x / mask");
  L2HMC_REQUIRE(a->L >= 1 && a->L <= kMaxL, "l2hmc_mog_mc_step: trajectory length must be in [1, %d]", kMaxL);
  L2HMC_REQUIRE(a->temperature > 0.f, "l2hmc_mog_mc_step: temperature must be positive");
  L2HMC_REQUIRE(a->v_f || !a->v_b, "l2hmc_mog_mc_step: v_b given without v_f");
  p.a = *a;
  p.H = 16;
  if (!a->hmc) {
    L2HMC_REQUIRE(a->xnet && a->vnet, "l2hmc_mog_mc_step: networks missing (hmc = 0)");
    for (int i = 0; i < 2; ++i) {
      const l2hmc_net_weights* w = i == 0 ? a->xnet : a->vnet;
      L2HMC_REQUIRE(w->K == p.k && w->D == p.k && w->H >= 1 && w->H <= 64 && w->H == a->xnet->H,
                    "l2hmc_mog_mc_step: nets must have K = D = k and equal H <= 64 (got K=%d D=%d H=%d)", w->K, w->D,
                    w->H);
      L2HMC_REQUIRE(w->w_in0 && w->b_in0 && w->w_in1 && w->b_in1 && w->w_t && w->b_t && w->w_h && w->b_h && w->w_s &&
                        w->b_s && w->w_tr && w->b_tr && w->w_q && w->b_q && w->coeff_s && w->coeff_q,
                    "l2hmc_mog_mc_step: null weight pointer");
      p.net[i] = *w;
    }
    p.H = a->xnet->H;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.k <= 2) return launch_mog_h<2>(p, st);
  if (p.k <= 4) return launch_mog_h<4>(p, st);
  return launch_mog_h<8>(p, st);
}

// ------------------------------------------------------------------ MoG training loss (value)
namespace l2hmc {
namespace {
constexpr int kLossBlocks = 1024;
// stage 1: per block, in fixed order, the four sums  sum 1/v1, sum 1/v2, sum v1, sum v2  with
// v1 = |x - Lx|^2 * px + 1e-4,  v2 = |z - Lz|^2 * pz + 1e-4  (mog_model.py:344-347); v1, v2 optionally stored
__global__ void __launch_bounds__(256) mog_loss_stage1_kernel(const float* __restrict__ x, const float* __restrict__ Lx,
                                                              const float* __restrict__ px, const float* __restrict__ z,
                                                              const float* __restrict__ Lz, const float* __restrict__ pz,
                                                              long long N, int k, float* v1_out, float* v2_out,
                                                              float* partial /* [gridDim.x][4] */) {
  __shared__ float sm[4][8];
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (long long n = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; n < N; n += static_cast<long long>(gridDim.x) * 256) {
    float d1 = 0.f, d2 = 0.f;
    for (int j = 0; j < k; ++j) {
      const float a = __fsub_rn(x[n * k + j], Lx[n * k + j]), b = __fsub_rn(z[n * k + j], Lz[n * k + j]);
      d1 = __fadd_rn(d1, __fmul_rn(a, a));
      d2 = __fadd_rn(d2, __fmul_rn(b, b));
    }
    const float v1 = __fadd_rn(__fmul_rn(d1, px[n]), 1e-4f), v2 = __fadd_rn(__fmul_rn(d2, pz[n]), 1e-4f);
    if (v1_out) v1_out[n] = v1;
    if (v2_out) v2_out[n] = v2;
    acc[0] += __fdiv_rn(1.f, v1), acc[1] += __fdiv_rn(1.f, v2), acc[2] += v1, acc[3] += v2;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], o);
    if (lane == 0) sm[q][warp] = acc[q];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sm[threadIdx.x][w];
    partial[blockIdx.x * 4 + threadIdx.x] = t;
  }
}
// stage 2: blocks in fixed order (double accumulation of <= 1024 partials), then
// loss = scale * (mean 1/v1 + mean 1/v2) + (-mean v1 - mean v2) / scale   (mog_model.py:350-355)
__global__ void mog_loss_stage2_kernel(const float* __restrict__ partial, int nb, long long N, float scale, float* loss) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double t[4] = {0., 0., 0., 0.};
  for (int b = 0; b < nb; ++b)
    for (int q = 0; q < 4; ++q) t[q] += static_cast<double>(partial[b * 4 + q]);
  const float m[4] = {static_cast<float>(t[0] / N), static_cast<float>(t[1] / N), static_cast<float>(t[2] / N),
                      static_cast<float>(t[3] / N)};
  float l = __fmul_rn(scale, __fadd_rn(m[0], m[1]));
  l = __fadd_rn(l, __fdiv_rn(__fsub_rn(-m[2], m[3]), scale));
  loss[0] = l;
}
}  // namespace
}  // namespace l2hmc

extern "C" size_t l2hmc_mog_loss_workspace_bytes(void) { return sizeof(float) * 4 * l2hmc::kLossBlocks; }

extern "C" int l2hmc_mog_loss(const float* x, const float* Lx, const float* px, const float* z, const float* Lz,
                              const float* pz, long long N, int k, float scale, float* v1, float* v2, float* loss,
                              void* workspace, size_t workspace_bytes, void* stream) {
  L2HMC_REQUIRE(x && Lx && px && z && Lz && pz && loss, "l2hmc_mog_loss: null pointer");
  L2HMC_REQUIRE(N > 0 && k >= 1 && scale != 0.f, "l2hmc_mog_loss: bad sizes N=%lld k=%d scale=%g", N, k, static_cast<double>(scale));
  L2HMC_REQUIRE(workspace && workspace_bytes >= l2hmc_mog_loss_workspace_bytes(), "l2hmc_mog_loss: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = static_cast<int>(std::min<long long>(kLossBlocks, (N + 255) / 256));
  float* partial = static_cast<float*>(workspace);
  mog_loss_stage1_kernel<<<nb, 256, 0, st>>>(x, Lx, px, z, Lz, pz, N, k, v1, v2, partial);
  mog_loss_stage2_kernel<<<1, 32, 0, st>>>(partial, nb, N, scale, loss);
  return check_launch("mog_loss_kernel", 2);
}
