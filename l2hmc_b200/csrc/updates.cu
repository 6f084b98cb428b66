// K2: masked augmented-leapfrog updates with in-register Jacobian log-det, plus the
// Metropolis-Hastings accept probability and direction/accept blends.
//
// Replaces (reference l2hmc/dynamics/gauge_dynamics.py):
//   :486-508 _update_momentum_forward     :537-561 _update_momentum_backward
//   :511-534 _update_position_forward     :565-590 _update_position_backward
//   :592-609 _compute_accept_prob         :683-689 kinetic_energy
//   :220-257 apply_transition's forward/accept mask blends
//
// Layout: one warp per chain, lanes stride the chain in float4 (scalar tail when D % 4 != 0),
// the per-chain log-det is a lane partial closed by a shuffle tree and ACCUMULATED into
// logdet[n] (one writer per chain: deterministic).  Every product/sum is rounded separately in
// the reference's association order (compiled with -fmad=false); exp() is expf (<= 1 ulp).
// Traffic per link: 5 reads + 1 write = 24 B with S,T,Q, 12 B without.
#include <algorithm>

#include "common.cuh"
#include "update_math.cuh"

namespace l2hmc {
namespace {

constexpr int kWarps = 8;

template <int DIR, bool NETS>
__global__ void __launch_bounds__(kWarps * 32) update_momentum_kernel(
    const float* v, const float* __restrict__ g, const float* __restrict__ S,
    const float* __restrict__ Tr, const float* __restrict__ Q, float eps, int N, int D, float* v_out, float* logdet,
    int32_t* flag) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float heps = __fmul_rn(0.5f, eps);
  const bool vec = (D & 3) == 0;
  for (int n = blockIdx.x * kWarps + warp; n < N; n += gridDim.x * kWarps) {
    const size_t base = static_cast<size_t>(n) * D;
    float ld = 0.f;
    bool bad = false;
    if (vec) {
      for (int i = lane * 4; i < D; i += 128) {
        const float4 vv = *reinterpret_cast<const float4*>(v + base + i);
        const float4 gg = *reinterpret_cast<const float4*>(g + base + i);
        float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f), t4 = s4, q4 = s4;
        if (NETS) {
          s4 = *reinterpret_cast<const float4*>(S + base + i);
          t4 = *reinterpret_cast<const float4*>(Tr + base + i);
          q4 = *reinterpret_cast<const float4*>(Q + base + i);
        }
        float4 o;
        o.x = momentum_elem<DIR>(vv.x, gg.x, s4.x, t4.x, q4.x, eps, heps, ld, bad);
        o.y = momentum_elem<DIR>(vv.y, gg.y, s4.y, t4.y, q4.y, eps, heps, ld, bad);
        o.z = momentum_elem<DIR>(vv.z, gg.z, s4.z, t4.z, q4.z, eps, heps, ld, bad);
        o.w = momentum_elem<DIR>(vv.w, gg.w, s4.w, t4.w, q4.w, eps, heps, ld, bad);
        *reinterpret_cast<float4*>(v_out + base + i) = o;
      }
    } else {
      for (int i = lane; i < D; i += 32) {
        const float s = NETS ? S[base + i] : 0.f, t = NETS ? Tr[base + i] : 0.f, q = NETS ? Q[base + i] : 0.f;
        v_out[base + i] = momentum_elem<DIR>(v[base + i], g[base + i], s, t, q, eps, heps, ld, bad);
      }
    }
    ld = warp_sum(ld);
    if (lane == 0 && logdet != nullptr) logdet[n] = __fadd_rn(logdet[n], ld);
    flag_nonfinite(flag, bad);
  }
}

// ---- position ---------------------------------------------------------------------------
template <int DIR, bool NETS>
__global__ void __launch_bounds__(kWarps * 32) update_position_kernel(
    const float* x, const float* __restrict__ v, const float* __restrict__ S,
    const float* __restrict__ Tr, const float* __restrict__ Q, const float* __restrict__ mask, int invert, float eps,
    int N, int D, float* x_out, float* logdet, int32_t* flag) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool vec = (D & 3) == 0;
  for (int n = blockIdx.x * kWarps + warp; n < N; n += gridDim.x * kWarps) {
    const size_t base = static_cast<size_t>(n) * D;
    float ld = 0.f;
    bool bad = false;
    if (vec) {
      for (int i = lane * 4; i < D; i += 128) {
        const float4 xx = *reinterpret_cast<const float4*>(x + base + i);
        const float4 vv = *reinterpret_cast<const float4*>(v + base + i);
        float4 mm = *reinterpret_cast<const float4*>(mask + i);
        if (invert) mm = make_float4(1.f - mm.x, 1.f - mm.y, 1.f - mm.z, 1.f - mm.w);
        float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f), t4 = s4, q4 = s4;
        if (NETS) {
          s4 = *reinterpret_cast<const float4*>(S + base + i);
          t4 = *reinterpret_cast<const float4*>(Tr + base + i);
          q4 = *reinterpret_cast<const float4*>(Q + base + i);
        }
        float4 o;
        o.x = position_elem<DIR>(xx.x, vv.x, s4.x, t4.x, q4.x, mm.x, eps, ld, bad);
        o.y = position_elem<DIR>(xx.y, vv.y, s4.y, t4.y, q4.y, mm.y, eps, ld, bad);
        o.z = position_elem<DIR>(xx.z, vv.z, s4.z, t4.z, q4.z, mm.z, eps, ld, bad);
        o.w = position_elem<DIR>(xx.w, vv.w, s4.w, t4.w, q4.w, mm.w, eps, ld, bad);
        *reinterpret_cast<float4*>(x_out + base + i) = o;
      }
    } else {
      for (int i = lane; i < D; i += 32) {
        const float s = NETS ? S[base + i] : 0.f, t = NETS ? Tr[base + i] : 0.f, q = NETS ? Q[base + i] : 0.f;
        const float m = invert ? 1.f - mask[i] : mask[i];
        x_out[base + i] = position_elem<DIR>(x[base + i], v[base + i], s, t, q, m, eps, ld, bad);
      }
    }
    ld = warp_sum(ld);
    if (lane == 0 && logdet != nullptr) logdet[n] = __fadd_rn(logdet[n], ld);
    flag_nonfinite(flag, bad);
  }
}

__global__ void __launch_bounds__(256) mask_mul_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                                        int invert, size_t total, int D, float* out) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float m = mask[i % D];
    out[i] = __fmul_rn(invert ? 1.f - m : m, x[i]);
  }
}

__global__ void __launch_bounds__(kWarps * 32) kinetic_kernel(const float* __restrict__ v, int N, int D, float* ke) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int n = blockIdx.x * kWarps + warp; n < N; n += gridDim.x * kWarps) {
    const size_t base = static_cast<size_t>(n) * D;
    float acc = 0.f;
    for (int i = lane; i < D; i += 32) {
      const float a = v[base + i];
      acc = __fadd_rn(acc, __fmul_rn(a, a));
    }
    acc = warp_sum(acc);
    if (lane == 0) ke[n] = __fmul_rn(0.5f, acc);
  }
}

__global__ void accept_prob_kernel(const float* __restrict__ h_old, const float* __restrict__ h_new,
                                   const float* __restrict__ sld, int N, float* p, int32_t* flag) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float d = __fsub_rn(h_old[n], h_new[n]);
  if (sld) d = __fadd_rn(d, sld[n]);
  float prob = expf(fminf(d, 0.f));
  if (isnan(d)) {  // tf.minimum propagates NaN -> exp -> check_numerics / is_finite -> 0
    prob = 0.f;
    if (flag) atomicOr(flag, 1);
  }
  p[n] = prob;
}

template <int VEC> struct RawVec { using type = float; };
template <> struct RawVec<4> { using type = float4; };

// One warp per chain.  The output arrays may alias the inputs, so the compiler cannot move the next
// iteration's loads above this iteration's stores: UNR vectors per stream are therefore loaded
// explicitly before any of them is consumed (bytes in flight = what keeps HBM busy).
// VEC = 4: D % 4 == 0 and 16-byte aligned arrays (every lattice here).  FULL: blends of both
// directions and / or the momenta are requested too (more streams, fewer vectors each).
template <int VEC, bool FULL>
__global__ void __launch_bounds__(kWarps * 32) mh_select_kernel(
    const float* __restrict__ x, const float* __restrict__ x_f, const float* __restrict__ x_b,
    const float* __restrict__ v_f, const float* __restrict__ v_b, const float* __restrict__ p_f,
    const float* __restrict__ p_b, const float* __restrict__ u_dir, const float* __restrict__ u_acc, int N, int D,
    int wrap, float* x_post, float* v_post, float* p, float* x_out, uint8_t* accepted) {
  constexpr int UNR = FULL ? 2 : 4;
  struct alignas(4 * VEC) Vec { float e[VEC]; };
  using Raw = typename RawVec<VEC>::type;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool both = FULL && x_b != nullptr;
  for (int n = blockIdx.x * kWarps + warp; n < N; n += gridDim.x * kWarps) {
    const size_t base = static_cast<size_t>(n) * D;
    auto ld = [&](const float* a, int i) {
      const Raw r = __ldcs(reinterpret_cast<const Raw*>(a + base + i));
      return *reinterpret_cast<const Vec*>(&r);
    };
    auto st = [&](float* a, int i, const Vec& w) {
      __stcs(reinterpret_cast<Raw*>(a + base + i), *reinterpret_cast<const Raw*>(&w));
    };
    float fm = 1.f, bm = 0.f, am = 1.f, rm = 0.f;
    for (int i0 = lane * VEC; i0 < D; i0 += 32 * VEC * UNR) {
      Vec xf[UNR], xb[UNR], x0[UNR], vf[UNR], vb[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int i = i0 + u * 32 * VEC;
        if (i < D) {
          xf[u] = ld(x_f, i);
          if (x_out) x0[u] = ld(x, i);
          if (both) xb[u] = ld(x_b, i);
          if (FULL && v_post) {
            vf[u] = ld(v_f, i);
            if (both) vb[u] = ld(v_b, i);
          }
        }
      }
      if (i0 == lane * VEC) {          // the chain's scalars, requested after the first vectors
        fm = (both && u_dir) ? (u_dir[n] > 0.5f ? 1.f : 0.f) : 1.f;
        bm = __fsub_rn(1.f, fm);
        float prob = p_f[n];
        if (both) prob = __fadd_rn(__fmul_rn(fm, prob), __fmul_rn(bm, p_b[n]));
        am = u_acc ? (prob > u_acc[n] ? 1.f : 0.f) : 1.f;
        rm = __fsub_rn(1.f, am);
        if (lane == 0) {
          if (p) p[n] = prob;
          if (accepted) accepted[n] = am != 0.f;
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int i = i0 + u * 32 * VEC;
        if (i >= D) continue;
        Vec xp = xf[u];
        if (both) {
#pragma unroll
          for (int j = 0; j < VEC; ++j) xp.e[j] = __fadd_rn(__fmul_rn(fm, xp.e[j]), __fmul_rn(bm, xb[u].e[j]));
        }
        if (x_post) st(x_post, i, xp);
        if (FULL && v_post) {
          Vec vp = vf[u];
          if (both) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) vp.e[j] = __fadd_rn(__fmul_rn(fm, vp.e[j]), __fmul_rn(bm, vb[u].e[j]));
          }
          st(v_post, i, vp);
        }
        if (x_out) {
          Vec o;
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            o.e[j] = __fadd_rn(__fmul_rn(am, xp.e[j]), __fmul_rn(rm, x0[u].e[j]));
            if (wrap) o.e[j] = wrap_2pi(o.e[j]);
          }
          st(x_out, i, o);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256) philox_draws_kernel(unsigned long long seed, unsigned long long mc_step,
                                                            const unsigned long long* mc_step_dev,
                                                            unsigned long long chain0, int N, int D, float* normals,
                                                            float* u_dir, float* u_acc) {
  if (mc_step_dev != nullptr) mc_step += *mc_step_dev;     // replayed CUDA graphs: the step counter lives on the device
  const int quads = (D + 3) / 4;
  const size_t total = static_cast<size_t>(N) * quads;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int n = static_cast<int>(i / quads), q = static_cast<int>(i % quads);
    if (normals) {
      float z[4];
      philox_normal4(seed, chain0 + n, mc_step, static_cast<uint32_t>(q), z);
      for (int j = 0; j < 4 && 4 * q + j < D; ++j) normals[static_cast<size_t>(n) * D + 4 * q + j] = z[j];
    }
    if (q == 0 && (u_dir || u_acc)) {
      float ud, ua;
      philox_uniform2(seed, chain0 + n, mc_step, ud, ua);
      if (u_dir) u_dir[n] = ud;
      if (u_acc) u_acc[n] = ua;
    }
  }
}

int chain_grid(int N) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return max(1, min((N + kWarps - 1) / kWarps, sms * 8));
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_update_momentum(const float* v, const float* g, const float* S, const float* Tr, const float* Q,
                                     float eps, int dir, int N, int D, float* v_out, float* logdet,
                                     int32_t* nonfinite_flag, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(v && g && v_out, "l2hmc_update_momentum: null pointer");
  L2HMC_REQUIRE(N >= 0 && D > 0, "l2hmc_update_momentum: bad sizes N=%d D=%d", N, D);
  L2HMC_REQUIRE(dir == 1 || dir == -1, "l2hmc_update_momentum: dir must be +1 or -1");
  const bool nets = S || Tr || Q;
  L2HMC_REQUIRE(!nets || (S && Tr && Q), "l2hmc_update_momentum: S, T, Q must be given together");
  if ((D & 3) == 0)
    L2HMC_REQUIRE(aligned16(v) && aligned16(g) && aligned16(v_out) && aligned16(S) && aligned16(Tr) && aligned16(Q),
                  "l2hmc_update_momentum: pointers must be 16-byte aligned");
  if (N == 0) return L2HMC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = chain_grid(N);
#define L2HMC_GO(DIRV, NETSV) \
  update_momentum_kernel<DIRV, NETSV><<<grid, kWarps * 32, 0, st>>>(v, g, S, Tr, Q, eps, N, D, v_out, logdet, nonfinite_flag)
  if (dir > 0) { if (nets) L2HMC_GO(1, true); else L2HMC_GO(1, false); }
  else { if (nets) L2HMC_GO(-1, true); else L2HMC_GO(-1, false); }
#undef L2HMC_GO
  return check_launch("update_momentum_kernel");
}

extern "C" int l2hmc_update_position(const float* x, const float* v, const float* S, const float* Tr, const float* Q,
                                     const float* mask, int invert_mask, float eps, int dir, int N, int D,
                                     float* x_out, float* logdet, int32_t* nonfinite_flag, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x && v && mask && x_out, "l2hmc_update_position: null pointer");
  L2HMC_REQUIRE(N >= 0 && D > 0, "l2hmc_update_position: bad sizes N=%d D=%d", N, D);
  L2HMC_REQUIRE(dir == 1 || dir == -1, "l2hmc_update_position: dir must be +1 or -1");
  const bool nets = S || Tr || Q;
  L2HMC_REQUIRE(!nets || (S && Tr && Q), "l2hmc_update_position: S, T, Q must be given together");
  if ((D & 3) == 0)
    L2HMC_REQUIRE(aligned16(x) && aligned16(v) && aligned16(x_out) && aligned16(mask) && aligned16(S) &&
                      aligned16(Tr) && aligned16(Q),
                  "l2hmc_update_position: pointers must be 16-byte aligned");
  if (N == 0) return L2HMC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = chain_grid(N);
#define L2HMC_GO(DIRV, NETSV)                                                                                   \
  update_position_kernel<DIRV, NETSV><<<grid, kWarps * 32, 0, st>>>(x, v, S, Tr, Q, mask, invert_mask, eps, N, D, \
                                                                    x_out, logdet, nonfinite_flag)
  if (dir > 0) { if (nets) L2HMC_GO(1, true); else L2HMC_GO(1, false); }
  else { if (nets) L2HMC_GO(-1, true); else L2HMC_GO(-1, false); }
#undef L2HMC_GO
  return check_launch("update_position_kernel");
}

extern "C" int l2hmc_mask_mul(const float* x, const float* mask, int invert_mask, int N, int D, float* out,
                              void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x && mask && out, "l2hmc_mask_mul: null pointer");
  L2HMC_REQUIRE(N >= 0 && D > 0, "l2hmc_mask_mul: bad sizes");
  if (N == 0) return L2HMC_OK;
  const size_t total = static_cast<size_t>(N) * D;
  const int grid = static_cast<int>(std::min<size_t>((total + 255) / 256, 148 * 16));
  mask_mul_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, mask, invert_mask, total, D, out);
  return check_launch("mask_mul_kernel");
}

extern "C" int l2hmc_kinetic_energy(const float* v, int N, int D, float* ke, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(v && ke, "l2hmc_kinetic_energy: null pointer");
  L2HMC_REQUIRE(N >= 0 && D > 0, "l2hmc_kinetic_energy: bad sizes");
  if (N == 0) return L2HMC_OK;
  kinetic_kernel<<<chain_grid(N), kWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(v, N, D, ke);
  return check_launch("kinetic_kernel");
}

extern "C" int l2hmc_accept_prob(const float* h_old, const float* h_new, const float* sumlogdet, int N, float* p,
                                 int32_t* nonfinite_flag, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(h_old && h_new && p, "l2hmc_accept_prob: null pointer");
  L2HMC_REQUIRE(N >= 0, "l2hmc_accept_prob: bad N");
  if (N == 0) return L2HMC_OK;
  accept_prob_kernel<<<(N + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(h_old, h_new, sumlogdet, N, p,
                                                                                      nonfinite_flag);
  return check_launch("accept_prob_kernel");
}

extern "C" int l2hmc_mh_select(const float* x, const float* x_f, const float* x_b, const float* v_f, const float* v_b,
                               const float* p_f, const float* p_b, const float* u_dir, const float* u_acc, int N,
                               int D, int wrap_2pi, float* x_post, float* v_post, float* p, float* x_out,
                               uint8_t* accepted, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x && x_f && p_f, "l2hmc_mh_select: null pointer");
  L2HMC_REQUIRE(!v_post || v_f, "l2hmc_mh_select: v_post requested without v_f");
  L2HMC_REQUIRE(!x_b || (p_b && (!v_post || v_b)), "l2hmc_mh_select: backward arrays must be given together");
  L2HMC_REQUIRE(N >= 0 && D > 0, "l2hmc_mh_select: bad sizes");
  if (N == 0) return L2HMC_OK;
  const auto misaligned = [](const void* a) { return (reinterpret_cast<uintptr_t>(a) & 15u) != 0; };
  const bool vec4 = D % 4 == 0 && !(misaligned(x) || misaligned(x_f) || misaligned(x_b) || misaligned(v_f) ||
                                    misaligned(v_b) || misaligned(x_post) || misaligned(v_post) || misaligned(x_out));
  const bool full = x_b != nullptr || v_post != nullptr;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define L2HMC_MH(V, F)                                                                                          \
  mh_select_kernel<V, F><<<chain_grid(N), kWarps * 32, 0, st>>>(x, x_f, x_b, v_f, v_b, p_f, p_b, u_dir, u_acc, N, D, \
                                                                wrap_2pi, x_post, v_post, p, x_out, accepted)
  if (vec4 && full) L2HMC_MH(4, true);
  else if (vec4) L2HMC_MH(4, false);
  else if (full) L2HMC_MH(1, true);
  else L2HMC_MH(1, false);
#undef L2HMC_MH
  return check_launch("mh_select_kernel");
}

// one thread: *ctr += inc (the Philox step counter of a captured MC step, advanced inside the graph)
__global__ void counter_add_kernel(unsigned long long* ctr, unsigned long long inc) { *ctr += inc; }

extern "C" int l2hmc_counter_add(uint64_t* counter, uint64_t inc, void* stream) {
  L2HMC_REQUIRE(counter, "l2hmc_counter_add: null counter");
  counter_add_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned long long*>(counter), inc);
  return check_launch("counter_add_kernel");
}

extern "C" int l2hmc_philox_draws(uint64_t seed, uint64_t mc_step, uint64_t chain0, int N, int D, float* normals,
                                  float* u_dir, float* u_acc, void* stream) {
  return l2hmc_philox_draws_dev(seed, mc_step, nullptr, chain0, N, D, normals, u_dir, u_acc, stream);
}

extern "C" int l2hmc_philox_draws_dev(uint64_t seed, uint64_t mc_step, const uint64_t* mc_step_dev, uint64_t chain0, int N,
                                      int D, float* normals, float* u_dir, float* u_acc, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(N >= 0 && D > 0, "l2hmc_philox_draws: bad sizes");
  L2HMC_REQUIRE(normals || u_dir || u_acc, "l2hmc_philox_draws: nothing requested");
  if (N == 0) return L2HMC_OK;
  const size_t total = static_cast<size_t>(N) * ((D + 3) / 4);
  const int grid = static_cast<int>(std::min<size_t>((total + 255) / 256, 148 * 16));
  philox_draws_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      seed, mc_step, reinterpret_cast<const unsigned long long*>(mc_step_dev), chain0, N, D, normals, u_dir, u_acc);
  return check_launch("philox_draws_kernel");
}
