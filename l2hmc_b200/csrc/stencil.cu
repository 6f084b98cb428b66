// K1 / K4: plaquette stencil for any T x X lattice -- Wilson action, force, plaquette and
// topological-charge observables in one pass over the links.
//
// Replaces (reference l2hmc/...):
//   lattice/lattice.py:337-362      GaugeLattice.total_action
//   lattice/lattice.py:285-313      GaugeLattice.calc_plaq_observables
//   dynamics/gauge_dynamics.py:698-709  GaugeDynamics.grad_potential (tf.gradients of beta*S)
//   gauge_model.py:659-716          _calc_plaq_sums/_calc_total_actions/_calc_avg_plaqs/_calc_top_charges
//
// One warp owns one chain at a time: the chain's D links are staged once through shared
// memory (coalesced float2 loads; each link is used by three plaquettes), beta*sin P goes to a
// second per-warp buffer for the force's backward differences, and the per-chain sums are
// lane-strided partials closed by a shuffle tree (deterministic).  Only __syncwarp is needed.
// HBM traffic: 4 B/link read (+ 4 B/link written when the force is requested).
#include "common.cuh"
#include "fast_sin.cuh"

namespace l2hmc {
namespace {

// cos P and project_angle(P) from ONE range reduction.  With j = rint(P / 2pi) and r = P - j 2pi (Cody-Waite,
// |r| <= pi), the reference's q = floor((P + pi) / 2pi) equals j whenever r is not within 1e-4 of +-pi (the
// quotient's fractional part is then in (1.5e-5, 1 - 1.5e-5), far outside any rounding), and
// proj = P - 2pi q is formed with the reference's own two roundings -- bit-identical to project_angle().
// Near the branch cut (probability ~3e-5) and for |P| > 4e4 the exact division is taken.
__device__ __forceinline__ void cos_and_project(float P, float* cs, float* proj) {
  if (__builtin_expect(fabsf(P) > 40000.f, 0)) {
    *cs = cosf(P);
    *proj = project_angle(P);
    return;
  }
  const float magic = 12582912.0f;
  const float j = __fsub_rn(fmaf(P, 0.15915494309189535f, magic), magic);
  float r = fmaf(j, -6.28125f, P);
  r = fmaf(j, -0.0019350051879882812f, r);
  r = fmaf(j, -3.019916050561733e-07f, r);
  const float a = fabsf(r);
  const bool fold = a > 1.57079637050628662f;
  const float folded = __fadd_rn(__fsub_rn(3.14159274101257324f, a), -8.742277657347586e-08f);
  const float c = cos_poly(fold ? folded : a);
  *cs = fold ? -c : c;
  *proj = a < 3.1415f ? __fsub_rn(P, __fmul_rn(kTwoPi, j)) : project_angle(P);
}

struct StencilParams {
  const float* x;
  float beta;
  int N, T, X;
  float* action;
  float* force;
  float* avg_plaq;
  float* q_float;
  int32_t* q_int;
  const float* v;   // with `hamil`: kinetic term
  float* hamil;     // beta*S + 0.5*sum v^2
};

__global__ void __launch_bounds__(256) stencil_kernel(const StencilParams p, int warps_per_block) {
  extern __shared__ __align__(16) float s_all[];
  const int T = p.T, X = p.X, SITES = T * X, D = 2 * SITES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* sx = s_all + static_cast<size_t>(warp) * ((D + SITES + 1) & ~1);  // keep float2 alignment
  float* sb = sx + D;
  const bool want_force = p.force != nullptr;
  const bool want_cos = p.action != nullptr || p.avg_plaq != nullptr || p.hamil != nullptr;
  const bool want_q = p.q_float != nullptr || p.q_int != nullptr;

  for (int n = blockIdx.x * warps_per_block + warp; n < p.N; n += gridDim.x * warps_per_block) {
    const float2* gx = reinterpret_cast<const float2*>(p.x + static_cast<size_t>(n) * D);
    for (int s = lane; s < SITES; s += 32) reinterpret_cast<float2*>(sx)[s] = gx[s];
    __syncwarp();
    float act = 0.f, csum = 0.f, qsum = 0.f;
    for (int s = lane; s < SITES; s += 32) {
      const int t = s / X, xx = s - t * X;
      const int sr = t * X + (xx + 1 == X ? 0 : xx + 1);   // (t, x+1)
      const int sd = (t + 1 == T ? 0 : t + 1) * X + xx;    // (t+1, x)
      const float P =
          __fadd_rn(__fsub_rn(__fsub_rn(sx[2 * s], sx[2 * s + 1]), sx[2 * sr]), sx[2 * sd + 1]);
      if (want_force && want_cos) {
        float sn, cs;
        fast_sincos(P, &sn, &cs);
        sb[s] = __fmul_rn(p.beta, sn);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
      } else if (want_force) {
        sb[s] = __fmul_rn(p.beta, fast_sin(P));
      } else if (want_cos && want_q) {
        float cs, pr;
        cos_and_project(P, &cs, &pr);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
        qsum = __fadd_rn(qsum, pr);
      } else if (want_cos) {
        const float cs = fast_cos(P);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
      }
      if (want_q && !(want_cos && !want_force)) qsum = __fadd_rn(qsum, project_angle(P));
    }
    if (want_force) {
      __syncwarp();
      float2* gf = reinterpret_cast<float2*>(p.force + static_cast<size_t>(n) * D);
      for (int s = lane; s < SITES; s += 32) {
        const int t = s / X, xx = s - t * X;
        const int sl = t * X + (xx == 0 ? X - 1 : xx - 1);  // (t, x-1)
        const int su = (t == 0 ? T - 1 : t - 1) * X + xx;   // (t-1, x)
        const float b = sb[s];
        gf[s] = make_float2(__fsub_rn(b, sb[sl]), __fsub_rn(sb[su], b));
      }
    }
    if (want_cos) {
      act = warp_sum(act);
      csum = warp_sum(csum);
    }
    if (want_q) qsum = warp_sum(qsum);
    if (p.hamil) {
      const float2* gv = reinterpret_cast<const float2*>(p.v + static_cast<size_t>(n) * D);
      float ke = 0.f;
      for (int s = lane; s < SITES; s += 32) {
        const float2 a = gv[s];
        ke = __fadd_rn(ke, __fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y)));
      }
      ke = warp_sum(ke);
      if (lane == 0) p.hamil[n] = __fadd_rn(__fmul_rn(p.beta, act), __fmul_rn(0.5f, ke));
    }
    if (lane == 0) {
      if (p.action) p.action[n] = act;
      if (p.avg_plaq) p.avg_plaq[n] = __fdiv_rn(csum, static_cast<float>(SITES));
      const float qf = __fdiv_rn(qsum, kTwoPi);
      if (p.q_float) p.q_float[n] = qf;
      if (p.q_int) p.q_int[n] = static_cast<int32_t>(floorf(__fadd_rn(0.1f, qf)));
    }
    __syncwarp();
  }
}


// Register-resident variant for X in {8, 16, 32} and T*X = 32*I sites: a warp holds one chain in
// registers (lane l owns sites l, l+32, ...: row i*(32/X) + l/X, column l%X), all 2*I float2 loads of
// a chain are in flight at once, neighbours come from warp shuffles (x +- 1) and the adjacent
// register (t +- 1; a second shuffle when a warp pass spans several rows).  No shared memory, so
// occupancy is set by registers alone.  Summation order is the generic kernel's (lane-strided
// partials, then the shuffle tree), hence bit-identical results.
template <int X, int I, bool FORCE, bool COS, bool QSUM, bool HAMIL>
__global__ void __launch_bounds__(256) stencil_reg_kernel(const StencilParams p) {
  constexpr int SITES = 32 * I, D = 2 * SITES;
  constexpr unsigned kFull = 0xffffffffu;
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int warps = (blockDim.x >> 5) * gridDim.x;
  const int right = (lane & ~(X - 1)) | ((lane + 1) & (X - 1));
  const int left = (lane & ~(X - 1)) | ((lane - 1) & (X - 1));
  const int down = (lane + X) & 31, up = (lane - X) & 31;
  const bool down_wraps = lane + X >= 32, up_wraps = lane < X;   // neighbour row lives in register i+1 / i-1
  for (int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps) {
    const float2* gx = reinterpret_cast<const float2*>(p.x + static_cast<size_t>(n) * D);
    float2 u[I];
#pragma unroll
    for (int i = 0; i < I; ++i) u[i] = __ldcs(gx + lane + 32 * i);
    float ke = 0.f;
    if (HAMIL) {
      const float2* gv = reinterpret_cast<const float2*>(p.v + static_cast<size_t>(n) * D);
      float2 w[I];
#pragma unroll
      for (int i = 0; i < I; ++i) w[i] = __ldcs(gv + lane + 32 * i);
#pragma unroll
      for (int i = 0; i < I; ++i)
        ke = __fadd_rn(ke, __fadd_rn(__fmul_rn(w[i].x, w[i].x), __fmul_rn(w[i].y, w[i].y)));
    }
    float b[FORCE ? I : 1];
    float act = 0.f, csum = 0.f, qsum = 0.f;
#pragma unroll
    for (int i = 0; i < I; ++i) {
      const int in = (i + 1 == I) ? 0 : i + 1;
      const float xr = __shfl_sync(kFull, u[i].x, right);              // x0 at (t, x+1)
      float yd;                                                        // x1 at (t+1, x)
      if (X == 32) {
        yd = u[in].y;
      } else {
        const float same = __shfl_sync(kFull, u[i].y, down), next = __shfl_sync(kFull, u[in].y, down);
        yd = down_wraps ? next : same;
      }
      const float P = __fadd_rn(__fsub_rn(__fsub_rn(u[i].x, u[i].y), xr), yd);
      if (FORCE && COS) {
        float sn, cs;
        fast_sincos(P, &sn, &cs);
        b[i] = __fmul_rn(p.beta, sn);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
      } else if (FORCE) {
        b[i] = __fmul_rn(p.beta, fast_sin(P));
      } else if (COS && QSUM) {
        float cs, pr;
        cos_and_project(P, &cs, &pr);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
        qsum = __fadd_rn(qsum, pr);
      } else if (COS) {
        const float cs = fast_cos(P);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
      }
      if (QSUM && !(COS && !FORCE)) qsum = __fadd_rn(qsum, project_angle(P));
    }
    if (FORCE) {
      float2* gf = reinterpret_cast<float2*>(p.force + static_cast<size_t>(n) * D);
#pragma unroll
      for (int i = 0; i < I; ++i) {
        const int ip = (i == 0) ? I - 1 : i - 1;
        const float bl = __shfl_sync(kFull, b[i], left);               // b at (t, x-1)
        float bu;                                                      // b at (t-1, x)
        if (X == 32) {
          bu = b[ip];
        } else {
          const float same = __shfl_sync(kFull, b[i], up), prev = __shfl_sync(kFull, b[ip], up);
          bu = up_wraps ? prev : same;
        }
        __stcs(gf + lane + 32 * i, make_float2(__fsub_rn(b[i], bl), __fsub_rn(bu, b[i])));
      }
    }
    if (COS) {
      act = warp_sum(act);
      csum = warp_sum(csum);
    }
    if (QSUM) qsum = warp_sum(qsum);
    if (HAMIL) {
      ke = warp_sum(ke);
      if (lane == 0) p.hamil[n] = __fadd_rn(__fmul_rn(p.beta, act), __fmul_rn(0.5f, ke));
    }
    if (lane == 0) {
      if (p.action) p.action[n] = act;
      if (p.avg_plaq) p.avg_plaq[n] = __fdiv_rn(csum, static_cast<float>(SITES));
      if (QSUM) {
        const float qf = __fdiv_rn(qsum, kTwoPi);
        if (p.q_float) p.q_float[n] = qf;
        if (p.q_int) p.q_int[n] = static_cast<int32_t>(floorf(__fadd_rn(0.1f, qf)));
      }
    }
  }
}

// Reduce-only variant (action / plaquette / charge, no force): the chain streams through a ring of PD row
// registers instead of living in registers whole -- ~40 registers instead of ~90, three times the resident
// warps, and PD - 1 loads in flight per lane all the time.  Same lane -> site map, same summation order and
// therefore the same bits as stencil_reg_kernel.
template <int X, int I, bool COS, bool QSUM>
__global__ void __launch_bounds__(256) stencil_roll_kernel(const StencilParams p) {
  constexpr int SITES = 32 * I, D = 2 * SITES, PD = I < 8 ? I : 8;
  static_assert(I % PD == 0, "row count must be a multiple of the ring depth");
  constexpr unsigned kFull = 0xffffffffu;
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int warps = (blockDim.x >> 5) * gridDim.x;
  const int right = (lane & ~(X - 1)) | ((lane + 1) & (X - 1));
  const int down = (lane + X) & 31;
  const bool down_wraps = lane + X >= 32;
  for (int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps) {
    const float2* gx = reinterpret_cast<const float2*>(p.x + static_cast<size_t>(n) * D) + lane;
    float2 ring[PD];
#pragma unroll
    for (int k = 0; k < PD; ++k) ring[k] = __ldcs(gx + 32 * k);
    const float2 first = ring[0];
    float act = 0.f, csum = 0.f, qsum = 0.f;
#pragma unroll 1
    for (int i0 = 0; i0 < I; i0 += PD) {
#pragma unroll
      for (int k = 0; k < PD; ++k) {
        const int i = i0 + k;
        const float2 cur = ring[k];
        if (i + PD < I) ring[k] = __ldcs(gx + 32 * (i + PD));          // row i + PD takes the slot just consumed
        const float2 nxt = (i + 1 == I) ? first : ring[(k + 1) % PD];  // row i + 1 (slot k + 1, or the reloaded slot 0)
        const float xr = __shfl_sync(kFull, cur.x, right);
        float yd;
        if (X == 32) {
          yd = nxt.y;
        } else {
          const float same = __shfl_sync(kFull, cur.y, down), next = __shfl_sync(kFull, nxt.y, down);
          yd = down_wraps ? next : same;
        }
        const float P = __fadd_rn(__fsub_rn(__fsub_rn(cur.x, cur.y), xr), yd);
        if (COS && QSUM) {
          float cs, pr;
          cos_and_project(P, &cs, &pr);
          act = __fadd_rn(act, __fsub_rn(1.f, cs));
          csum = __fadd_rn(csum, cs);
          qsum = __fadd_rn(qsum, pr);
        } else if (COS) {
          const float cs = fast_cos(P);
          act = __fadd_rn(act, __fsub_rn(1.f, cs));
          csum = __fadd_rn(csum, cs);
        } else {
          qsum = __fadd_rn(qsum, project_angle(P));
        }
      }
    }
    if (COS) {
      act = warp_sum(act);
      csum = warp_sum(csum);
    }
    if (QSUM) qsum = warp_sum(qsum);
    if (lane == 0) {
      if (p.action) p.action[n] = act;
      if (p.avg_plaq) p.avg_plaq[n] = __fdiv_rn(csum, static_cast<float>(SITES));
      if (QSUM) {
        const float qf = __fdiv_rn(qsum, kTwoPi);
        if (p.q_float) p.q_float[n] = qf;
        if (p.q_int) p.q_int[n] = static_cast<int32_t>(floorf(__fadd_rn(0.1f, qf)));
      }
    }
  }
}

template <int X, int I>
bool launch_reg(const StencilParams& p, int sms, cudaStream_t stream) {
  const bool force = p.force != nullptr, hamil = p.hamil != nullptr;
  const bool cosv = p.action != nullptr || p.avg_plaq != nullptr || hamil;
  const bool q = p.q_float != nullptr || p.q_int != nullptr;
  const int blocks = min((p.N + 7) / 8, sms * 16);
  // the combinations the API can ask for: force [+ action]; action / plaq / Q observables; Hamiltonian
  if (force && cosv && !q && !hamil) launch_pdl(stencil_reg_kernel<X, I, true, true, false, false>, dim3(blocks), dim3(256), 0, stream, p);
  else if (force && !cosv && !q) launch_pdl(stencil_reg_kernel<X, I, true, false, false, false>, dim3(blocks), dim3(256), 0, stream, p);
  else if (!force && cosv && !q && !hamil) launch_pdl(stencil_roll_kernel<X, I, true, false>, dim3(blocks), dim3(256), 0, stream, p);
  else if (!force && cosv && q && !hamil) launch_pdl(stencil_roll_kernel<X, I, true, true>, dim3(blocks), dim3(256), 0, stream, p);
  else if (!force && !cosv && q) launch_pdl(stencil_roll_kernel<X, I, false, true>, dim3(blocks), dim3(256), 0, stream, p);
  else if (!force && hamil && !q) launch_pdl(stencil_reg_kernel<X, I, false, true, false, true>, dim3(blocks), dim3(256), 0, stream, p);
  else return false;
  return true;
}

int launch_stencil(const StencilParams& p, cudaStream_t stream) {
  if (p.N == 0) return L2HMC_OK;
  const size_t per_warp = sizeof(float) * ((3u * p.T * p.X + 1u) & ~1u);
  int dev = 0, max_smem = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  bool done = false;
  if (p.T == 32 && p.X == 32) done = launch_reg<32, 32>(p, sms, stream);
  else if (p.T == 16 && p.X == 16) done = launch_reg<16, 8>(p, sms, stream);
  else if (p.T == 8 && p.X == 8) done = launch_reg<8, 2>(p, sms, stream);
  if (done) return check_launch("stencil_reg_kernel");
  int warps = 8;
  while (warps > 1 && per_warp * warps > 96 * 1024) warps >>= 1;
  const size_t smem = per_warp * warps;
  if (smem > static_cast<size_t>(max_smem)) {
    set_error("stencil: a %dx%d chain (%zu B) does not fit shared memory (%d B)", p.T, p.X, smem, max_smem);
    return L2HMC_ERR_UNSUPPORTED;
  }
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(stencil_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(stencil_kernel): %s", cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  const int blocks_needed = (p.N + warps - 1) / warps;
  const int grid = min(blocks_needed, sms * 16);
  stencil_kernel<<<grid, warps * 32, smem, stream>>>(p, warps);
  return check_launch("stencil_kernel");
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_u1_action_force(const float* x, float beta, int N, int T, int X, float* action, float* force,
                                     void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x, "l2hmc_u1_action_force: null x");
  L2HMC_REQUIRE(N >= 0 && T > 0 && X > 0, "l2hmc_u1_action_force: bad sizes N=%d T=%d X=%d", N, T, X);
  L2HMC_REQUIRE(action || force, "l2hmc_u1_action_force: nothing requested");
  StencilParams p{x, beta, N, T, X, action, force, nullptr, nullptr, nullptr, nullptr, nullptr};
  return launch_stencil(p, static_cast<cudaStream_t>(stream));
}

extern "C" int l2hmc_u1_observables(const float* x, int N, int T, int X, float* action, float* avg_plaq,
                                    float* q_float, int32_t* q_int, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x, "l2hmc_u1_observables: null x");
  L2HMC_REQUIRE(N >= 0 && T > 0 && X > 0, "l2hmc_u1_observables: bad sizes N=%d T=%d X=%d", N, T, X);
  L2HMC_REQUIRE(action || avg_plaq || q_float || q_int, "l2hmc_u1_observables: nothing requested");
  StencilParams p{x, 0.f, N, T, X, action, nullptr, avg_plaq, q_float, q_int, nullptr, nullptr};
  return launch_stencil(p, static_cast<cudaStream_t>(stream));
}

extern "C" int l2hmc_u1_hamiltonian(const float* x, const float* v, float beta, int N, int T, int X, float* h,
                                    void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x && v && h, "l2hmc_u1_hamiltonian: null pointer");
  L2HMC_REQUIRE(N >= 0 && T > 0 && X > 0, "l2hmc_u1_hamiltonian: bad sizes N=%d T=%d X=%d", N, T, X);
  StencilParams p{x, beta, N, T, X, nullptr, nullptr, nullptr, nullptr, nullptr, v, h};
  return launch_stencil(p, static_cast<cudaStream_t>(stream));
}
