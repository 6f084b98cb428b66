// Fused plain-HMC kernels for the 2D U(1) gauge model (S = T = Q = 0 path of L2HMC).
//
//   hmc_kernel<T, X, /*MC=*/false>  one leapfrog step per launch: x, v streamed once
//                                   (16 B / link-update) -- the HBM-roofline kernel.
//   hmc_kernel<T, X, /*MC=*/true>   a whole Monte-Carlo step per launch: momenta (injected or
//                                   Philox), M leapfrog steps with x, v resident on-chip, both
//                                   Hamiltonians, accept probability, direction + accept select,
//                                   mod-2pi wrap and the observables of the resulting state.
//
// Replaces GaugeDynamics.apply_transition / transition_kernel / _forward_lf / _backward_lf with
// the hmc stubs (reference l2hmc/dynamics/gauge_dynamics.py:102-108, 195-313, 412-483), the
// force tf.gradients(beta * total_action) (:698-709, lattice/lattice.py:337-362) and the
// observables of lattice/lattice.py:285-313.
//
// Structure (B200): persistent CTAs of 8 warps.  A tile is 2048 links = 8 KB of x (+ 8 KB of v)
// = 256/TPC whole chains.  Tiles move HBM <-> shared memory with 1-D bulk async copies (TMA
// engine): loads land in a 2-deep ring and complete on an mbarrier, issued TWO tiles ahead as
// soon as the ring slot has been read into registers; results are staged in shared memory and
// leave as bulk stores.  No thread ever waits on a global load, and no warp polls.
// Each thread owns a strip of four sites of one chain (8 links, contiguous) in registers as
// packed fp32x2 pairs; the plaquette needs the next row and column (a shared-memory plane of the
// mu=1 links / one shuffle), the force needs beta*sin P of the previous row and column (a second
// plane / one shuffle): per force evaluation two conflict-free float4 stores, two float4 loads,
// two shuffles and two named barriers.
//
// Arithmetic keeps the reference's fp32 op order (every product and sum rounded separately):
//   P  = ((U0 - U1) - U0[x+1]) + U1[t+1]            lattice.py:354-357
//   b  = beta * sin P;  g0 = b - b[x-1];  g1 = -b + b[t-1]   (adjoint of the roll graph)
//   v' = v - (eps/2) * g;   x' = x + eps * v'        gauge_dynamics.py:486-534 with S=T=Q=0
// and the backward step is the same with eps -> -eps (bit-identical to :537-590).  The math
// runs on FFMA2 / FADD2 / FMUL2 (two IEEE lanes per issue slot: with the trajectory on-chip the
// kernels are issue-bound, not HBM-bound); sin/cos are the branch-free fast_sin.cuh versions.
#include "common.cuh"
#include "fast_sin.cuh"

namespace l2hmc {

namespace {

constexpr int kThreads = 256;
constexpr int kTileFloats = 2048;  // links per tile (8 KB)

struct HmcParams {
  const float* x;
  const float* v_f;
  const float* v_b;
  const float* u_dir;
  const float* u_acc;
  const int8_t* dir;
  float beta, eps, one;
  int num_steps, default_dir, wrap, N, num_tiles;
  unsigned long long seed, mc_step, chain0;
  const unsigned long long* mc_step_dev;   // nullable: added to mc_step on the device (replayed CUDA graphs)
  float* x_post;
  float* v_post;
  float* accept_prob;
  float* x_out;
  uint8_t* accepted;
  float* h_old;
  float* h_new;
  float* obs_action;
  float* obs_plaq;
  int32_t* obs_charge;
  int32_t* flag;
  // shared-memory layout, in floats (filled by launch())
  int in_floats;   // one load-ring slot: x (+ v when momenta are injected)
  int off_out;     // start of the output staging area
  int off_xpost, off_vpost, off_xout;   // offsets inside the staging area, -1 = not requested
  int off_planes;  // two site planes (mu=1 links, beta*sin P)
  int off_red;     // per-warp partial sums
  int off_bar;     // two mbarriers
};

template <int T, int X>
struct Geo {
  static constexpr int SITES = T * X;
  static constexpr int D = 2 * SITES;
  static constexpr int TPC = SITES / 4;       // threads per chain
  static constexpr int CPB = kThreads / TPC;  // chains per tile
  static constexpr int XG = X / 4;            // threads per lattice row
  static_assert(X % 4 == 0, "strip of four sites per thread");
  static_assert(TPC >= 1 && TPC <= kThreads && kThreads % TPC == 0, "chain must fit one CTA");
  static_assert(XG >= 1 && XG <= 32 && (XG & (XG - 1)) == 0, "row neighbours by warp shuffle");
  static_assert(CPB * D == kTileFloats, "tile is always 8 KB");
};

// 128-bit shared-memory accesses by 32-bit shared address (no generic-pointer arithmetic in the loop)
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int T, int X, bool MC>
__global__ void __launch_bounds__(kThreads, MC ? 3 : 4) hmc_kernel(const HmcParams p) {
  using G = Geo<T, X>;
  constexpr int D = G::D, TPC = G::TPC, CPB = G::CPB, XG = G::XG;
  extern __shared__ __align__(128) float smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  float* s_red = smem + p.off_red;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const bool rng = MC && (p.v_f == nullptr);
  const bool per_chain_v = MC && !rng && p.v_b != nullptr;
  const unsigned long long mc_step = p.mc_step + ((MC && p.mc_step_dev != nullptr) ? *p.mc_step_dev : 0ull);

  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  // Warp 0 issues the bulk load of `tile` into ring slot `slot` (lane 0: links; momenta: one
  // copy, or one per chain from v_f / v_b according to the chain's direction).
  auto issue_load = [&](int tile, int slot) {
    const int chain_base = tile * CPB;
    const int nvalid = min(CPB, p.N - chain_base);
    const uint32_t bytes = static_cast<uint32_t>(nvalid) * D * 4u;
    float* sx = smem + slot * p.in_floats;
    const size_t goff = static_cast<size_t>(chain_base) * D;
    if (lane == 0) {
      mbar_arrive_expect_tx(&full[slot], rng ? bytes : 2u * bytes);
      bulk_g2s(sx, p.x + goff, bytes, &full[slot]);
      if (!rng && !per_chain_v) bulk_g2s(sx + kTileFloats, p.v_f + goff, bytes, &full[slot]);
    }
    __syncwarp();
    if (per_chain_v) {
      for (int cc = lane; cc < nvalid; cc += 32) {
        const bool fwd = p.u_dir ? (p.u_dir[chain_base + cc] > 0.5f) : (p.default_dir >= 0);
        const float* src = (fwd ? p.v_f : p.v_b) + goff + static_cast<size_t>(cc) * D;
        bulk_g2s(sx + kTileFloats + cc * D, src, D * 4u, &full[slot]);
      }
      __syncwarp();
    }
  };
  if (warp == 0) {
    issue_load(blockIdx.x, 0);
    if (blockIdx.x + gridDim.x < p.num_tiles) issue_load(blockIdx.x + gridDim.x, 1);
  }

  const int k = tid % TPC, c = tid / TPC;
  const int t = k / XG, cg = k % XG;
  const int x0 = 4 * cg;
  const int lane_right = (lane & ~(XG - 1)) | ((lane + 1) & (XG - 1));
  const int lane_left = (lane & ~(XG - 1)) | ((lane + XG - 1) & (XG - 1));
  const int off_own = c * D + 8 * k;  // own strip inside a tile buffer (floats)
  const uint32_t planes = smem_u32(smem + p.off_planes);
  const uint32_t a_u1_own = planes + 4u * (c * G::SITES + t * X + x0);               // mu=1 links, own sites
  const uint32_t a_u1_dn = planes + 4u * (c * G::SITES + ((t + 1) % T) * X + x0);    // ... of row t+1
  const uint32_t a_b_own = a_u1_own + 4u * (kTileFloats / 2);                        // beta*sin P, own sites
  const uint32_t a_b_up = planes + 4u * (kTileFloats / 2 + c * G::SITES + ((t + T - 1) % T) * X + x0);
  const float beta = p.beta;
  const float half_eps = __fmul_rn(0.5f, p.eps);
  const f32x2 one = pack2(p.one, p.one);  // runtime 1.0: keeps ptxas from contracting mul+add
  const f32x2 beta2 = pack2(beta, beta);

  // Thread state, packed fp32x2 in SoA order: X0a = (U0(s0), U0(s1)), X0b = (U0(s2), U0(s3)),
  // X1a/X1b the mu=1 links, same for momenta V* and force G* (s0..s3 = the thread's four sites).
  f32x2 X0a, X0b, X1a, X1b, V0a, V0b, V1a, V1b, G0a, G0b, G1a, G1b;

  auto publish_u1 = [&]() {
    sts128(a_u1_own, lo2(X1a), hi2(X1a), lo2(X1b), hi2(X1b));
    named_bar_sync(1, kThreads);
  };
  // Force from the links in registers, after publish_u1().  obs != nullptr additionally
  // accumulates {sum(1-cos P), sum cos P, sum proj(P)} (the trajectory's two ends).
  auto force = [&](float* obs) {
    const float4 dn = lds128(a_u1_dn);
    const float u0r = __shfl_sync(0xffffffffu, lo2(X0a), lane_right);
    // P = ((U0 - U1) - U0[x+1]) + U1[t+1]   (lattice.py:354-357), two plaquettes per instruction
    const f32x2 Pa = add2(sub2(sub2(X0a, X1a), pack2(hi2(X0a), lo2(X0b))), pack2(dn.x, dn.y));
    const f32x2 Pb = add2(sub2(sub2(X0b, X1b), pack2(hi2(X0b), u0r)), pack2(dn.z, dn.w));
    const float pmax = fmaxf(fmaxf(fabsf(lo2(Pa)), fabsf(hi2(Pa))), fmaxf(fabsf(lo2(Pb)), fabsf(hi2(Pb))));
    f32x2 Ba, Bb;
    if (obs != nullptr || !(pmax <= 40000.f)) {  // trajectory ends (need cos) or out-of-range angles
      const float P[4] = {lo2(Pa), hi2(Pa), lo2(Pb), hi2(Pb)};
      float b[4], act = 0.f, csum = 0.f, qsum = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float sn, cs, pr;
        sincos_and_project(P[j], &sn, &cs, &pr);    // one range reduction for all three (fast_sin.cuh)
        b[j] = __fmul_rn(beta, sn);
        act = __fadd_rn(act, __fsub_rn(1.f, cs));
        csum = __fadd_rn(csum, cs);
        qsum = __fadd_rn(qsum, pr);
      }
      if (obs != nullptr) obs[0] = act, obs[1] = csum, obs[2] = qsum;
      Ba = pack2(b[0], b[1]), Bb = pack2(b[2], b[3]);
    } else {
      Ba = mul2(beta2, fast_sin2(Pa));
      Bb = mul2(beta2, fast_sin2(Pb));
    }
    sts128(a_b_own, lo2(Ba), hi2(Ba), lo2(Bb), hi2(Bb));
    named_bar_sync(1, kThreads);
    const float4 bu = lds128(a_b_up);
    const float bl = __shfl_sync(0xffffffffu, hi2(Bb), lane_left);
    G0a = sub2(Ba, pack2(bl, lo2(Ba)));  // g0 = b - b[x-1]
    G0b = sub2(Bb, pack2(hi2(Ba), lo2(Bb)));
    G1a = sub2(pack2(bu.x, bu.y), Ba);   // g1 = -b + b[t-1]
    G1b = sub2(pack2(bu.z, bu.w), Bb);
  };

  constexpr int GW = TPC < 32 ? TPC : 32;  // lanes of a warp that belong to the same chain
  int it = 0;
  for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
    const int slot = it & 1;
    const float* sx = smem + slot * p.in_floats;
    const int chain = tile * CPB + c;
    const bool valid = chain < p.N;

    // the previous tile's bulk store must have drained the staging area before it is rewritten
    // (it is rewritten only after the many barriers below, which thread 0 reaches after this wait)
    if (tid == 0 && it > 0) bulk_wait_read_all();

    // per-chain direction / accept draw (global reads issued before the shared-memory wait)
    float sgn = 1.f, u_acc = 2.f;
    bool have_acc = false;
    if (MC) {
      if (rng) {
        float ud = 0.f, ua = 0.f;
        if ((lane & (GW - 1)) == 0)   // one Philox call per chain and warp, then broadcast
          philox_uniform2(p.seed, p.chain0 + static_cast<unsigned long long>(chain), mc_step, ud, ua);
        ud = __shfl_sync(0xffffffffu, ud, lane & ~(GW - 1));
        ua = __shfl_sync(0xffffffffu, ua, lane & ~(GW - 1));
        sgn = (p.u_dir ? (valid ? p.u_dir[chain] : 1.f) : ud) > 0.5f ? 1.f : -1.f;
        u_acc = p.u_acc ? (valid ? p.u_acc[chain] : 0.f) : ua;
        have_acc = true;
      } else {
        if (p.u_dir) sgn = (valid ? p.u_dir[chain] : 1.f) > 0.5f ? 1.f : -1.f;
        else sgn = p.default_dir >= 0 ? 1.f : -1.f;
        if (p.u_acc) {
          u_acc = valid ? p.u_acc[chain] : 0.f;
          have_acc = true;
        }
      }
    } else if (p.dir) {
      sgn = (valid ? p.dir[chain] : 1) >= 0 ? 1.f : -1.f;
    }
    // v' = v + (-s*eps/2)*g and x' = x + (s*eps)*v': the negated factor gives bit-identical results
    // to the reference's v - (eps/2)*g (and to its backward step for s = -1)
    const float eps_s = __fmul_rn(sgn, p.eps);
    const float nheps_s = -__fmul_rn(sgn, half_eps);
    const f32x2 eps2 = pack2(eps_s, eps_s), nheps2 = pack2(nheps_s, nheps_s);

    mbar_wait(&full[slot], (it >> 1) & 1);

    if (valid) {
      const float4 a0 = *reinterpret_cast<const float4*>(sx + off_own);
      const float4 a1 = *reinterpret_cast<const float4*>(sx + off_own + 4);
      X0a = pack2(a0.x, a0.z), X1a = pack2(a0.y, a0.w);
      X0b = pack2(a1.x, a1.z), X1b = pack2(a1.y, a1.w);
      if (rng) {
        float z[8];
        philox_normal4(p.seed, p.chain0 + static_cast<unsigned long long>(chain), mc_step, 2u * k, z);
        philox_normal4(p.seed, p.chain0 + static_cast<unsigned long long>(chain), mc_step, 2u * k + 1u, z + 4);
        V0a = pack2(z[0], z[2]), V1a = pack2(z[1], z[3]);
        V0b = pack2(z[4], z[6]), V1b = pack2(z[5], z[7]);
      } else {
        const float4 w0 = *reinterpret_cast<const float4*>(sx + kTileFloats + off_own);
        const float4 w1 = *reinterpret_cast<const float4*>(sx + kTileFloats + off_own + 4);
        V0a = pack2(w0.x, w0.z), V1a = pack2(w0.y, w0.w);
        V0b = pack2(w1.x, w1.z), V1b = pack2(w1.y, w1.w);
      }
    } else {  // tail tile: chains past N integrate zeros and are never stored
      X0a = X0b = X1a = X1b = V0a = V0b = V1a = V1b = 0ULL;
    }

    const f32x2 S0a = X0a, S0b = X0b, S1a = X1a, S1b = X1b;  // x at the start of the trajectory (reject)
    float obs0[3] = {0.f, 0.f, 0.f}, obs1[3] = {0.f, 0.f, 0.f};
    float ke0 = 0.f, ke1 = 0.f;
    auto kinetic = [&]() {  // sum v^2 in link order, every term and sum rounded separately
      const float v[8] = {lo2(V0a), lo2(V1a), hi2(V0a), hi2(V1a), lo2(V0b), lo2(V1b), hi2(V0b), hi2(V1b)};
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc = __fadd_rn(acc, __fmul_rn(v[j], v[j]));
      return acc;
    };
    if (MC) ke0 = kinetic();

    publish_u1();
    // every thread has now copied its strip out of the ring slot: refill it two tiles ahead
    if (warp == 0 && tile + 2 * static_cast<int>(gridDim.x) < p.num_tiles)
      issue_load(tile + 2 * static_cast<int>(gridDim.x), slot);
    force(MC ? obs0 : nullptr);
    const int nsteps = MC ? p.num_steps : 1;
    for (int step = 0; step < nsteps; ++step) {
      V0a = madd2_unfused(V0a, nheps2, G0a, one), V0b = madd2_unfused(V0b, nheps2, G0b, one);
      V1a = madd2_unfused(V1a, nheps2, G1a, one), V1b = madd2_unfused(V1b, nheps2, G1b, one);
      X0a = madd2_unfused(X0a, eps2, V0a, one), X0b = madd2_unfused(X0b, eps2, V0b, one);
      X1a = madd2_unfused(X1a, eps2, V1a, one), X1b = madd2_unfused(X1b, eps2, V1b, one);
      publish_u1();
      force((MC && step == nsteps - 1) ? obs1 : nullptr);
      V0a = madd2_unfused(V0a, nheps2, G0a, one), V0b = madd2_unfused(V0b, nheps2, G0b, one);
      V1a = madd2_unfused(V1a, nheps2, G1a, one), V1b = madd2_unfused(V1b, nheps2, G1b, one);
    }

    float* so = smem + p.off_out;
    if (p.off_xpost >= 0) {
      float* q = so + p.off_xpost + off_own;
      *reinterpret_cast<float4*>(q) = make_float4(lo2(X0a), lo2(X1a), hi2(X0a), hi2(X1a));
      *reinterpret_cast<float4*>(q + 4) = make_float4(lo2(X0b), lo2(X1b), hi2(X0b), hi2(X1b));
    }
    if (p.off_vpost >= 0) {
      float* q = so + p.off_vpost + off_own;
      *reinterpret_cast<float4*>(q) = make_float4(lo2(V0a), lo2(V1a), hi2(V0a), hi2(V1a));
      *reinterpret_cast<float4*>(q + 4) = make_float4(lo2(V0b), lo2(V1b), hi2(V0b), hi2(V1b));
    }

    if (MC) {
      ke1 = kinetic();
      // per-chain sums of {action, kinetic, sum cos, sum proj} at both ends of the trajectory
      float r[8] = {obs0[0], ke0, obs0[1], obs0[2], obs1[0], ke1, obs1[1], obs1[2]};
      if (TPC <= 32) {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = group_sum<GW>(r[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = warp_sum(r[j]);
        if (lane == 0) {
          *reinterpret_cast<float4*>(s_red + 8 * warp) = make_float4(r[0], r[1], r[2], r[3]);
          *reinterpret_cast<float4*>(s_red + 8 * warp + 4) = make_float4(r[4], r[5], r[6], r[7]);
        }
        named_bar_sync(1, kThreads);
        constexpr int WPC = TPC > 32 ? TPC / 32 : 1;
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = 0.f;
#pragma unroll
        for (int w = 0; w < WPC; ++w) {
          const float4 q0 = *reinterpret_cast<const float4*>(s_red + 8 * (c * WPC + w));
          const float4 q1 = *reinterpret_cast<const float4*>(s_red + 8 * (c * WPC + w) + 4);
          r[0] += q0.x, r[1] += q0.y, r[2] += q0.z, r[3] += q0.w;
          r[4] += q1.x, r[5] += q1.y, r[6] += q1.z, r[7] += q1.w;
        }
      }
      const float h0 = __fadd_rn(__fmul_rn(beta, r[0]), __fmul_rn(0.5f, r[1]));
      const float h1 = __fadd_rn(__fmul_rn(beta, r[4]), __fmul_rn(0.5f, r[5]));
      const float dh = __fsub_rn(h0, h1);
      float prob = expf(fminf(dh, 0.f));
      if (isnan(dh)) prob = 0.f;  // tf.minimum propagates NaN; is_finite -> 0 (gauge_dynamics.py:604-609)
      const bool acc = have_acc ? (prob > u_acc) : true;
      if (valid && k == 0) {
        if (p.accept_prob) p.accept_prob[chain] = prob;
        if (p.accepted) p.accepted[chain] = acc ? 1 : 0;
        if (p.h_old) p.h_old[chain] = h0;
        if (p.h_new) p.h_new[chain] = h1;
        if (isnan(dh) && p.flag) atomicOr(p.flag, 1);
        // observables of the state the chain ends the MC step in (lattice.py:300-311)
        const float act = acc ? r[4] : r[0], csum = acc ? r[6] : r[2], qsum = acc ? r[7] : r[3];
        if (p.obs_action) p.obs_action[chain] = act;
        if (p.obs_plaq) p.obs_plaq[chain] = __fdiv_rn(csum, static_cast<float>(G::SITES));
        if (p.obs_charge) p.obs_charge[chain] = static_cast<int32_t>(floorf(__fadd_rn(0.1f, __fdiv_rn(qsum, kTwoPi))));
      }
      if (p.off_xout >= 0) {
        const f32x2 O0a = acc ? X0a : S0a, O0b = acc ? X0b : S0b, O1a = acc ? X1a : S1a, O1b = acc ? X1b : S1b;
        float o[8] = {lo2(O0a), lo2(O1a), hi2(O0a), hi2(O1a), lo2(O0b), lo2(O1b), hi2(O0b), hi2(O1b)};
        if (p.wrap) {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = wrap_2pi(o[j]);
        }
        float* q = so + p.off_xout + off_own;
        *reinterpret_cast<float4*>(q) = make_float4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<float4*>(q + 4) = make_float4(o[4], o[5], o[6], o[7]);
      }
    }
    // staged results -> HBM: one bulk store per requested output, issued by thread 0
    fence_proxy_async_smem();
    named_bar_sync(1, kThreads);
    if (tid == 0) {
      const int chain_base = tile * CPB;
      const uint32_t bytes = static_cast<uint32_t>(min(CPB, p.N - chain_base)) * D * 4u;
      const size_t goff = static_cast<size_t>(chain_base) * D;
      if (p.off_xpost >= 0) bulk_s2g(p.x_post + goff, so + p.off_xpost, bytes);
      if (p.off_vpost >= 0) bulk_s2g(p.v_post + goff, so + p.off_vpost, bytes);
      if (p.off_xout >= 0) bulk_s2g(p.x_out + goff, so + p.off_xout, bytes);
      bulk_commit();
    }
  }
  if (tid == 0) bulk_wait_read_all();  // shared memory must outlive the last store's reads
}

template <int T, int X, bool MC>
int launch(const HmcParams& p_in, cudaStream_t stream) {
  HmcParams p = p_in;
  using G = Geo<T, X>;
  auto kern = hmc_kernel<T, X, MC>;
  // shared-memory layout (floats): [ring slot 0][ring slot 1][staging][planes][partials][mbarriers]
  const bool rng = MC && p.v_f == nullptr;
  p.in_floats = rng ? kTileFloats : 2 * kTileFloats;
  int nout = 0;
  p.off_xpost = p.x_post ? (nout++) * kTileFloats : -1;
  p.off_vpost = p.v_post ? (nout++) * kTileFloats : -1;
  p.off_xout = (MC && p.x_out) ? (nout++) * kTileFloats : -1;
  p.off_out = 2 * p.in_floats;
  p.off_planes = p.off_out + nout * kTileFloats;
  p.off_red = p.off_planes + kTileFloats;
  p.off_bar = p.off_red + 64;
  const size_t smem_bytes = sizeof(float) * static_cast<size_t>(p.off_bar) + 2 * sizeof(uint64_t);

  // per device (one process may drive several GPUs): SM count and the opted-in shared-memory size
  constexpr int kMaxDevices = 64;
  static int sm_counts[kMaxDevices] = {};
  static size_t smem_opts[kMaxDevices] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  const int slot = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
  if (sm_counts[slot] == 0 || slot != dev)
    cudaDeviceGetAttribute(&sm_counts[slot], cudaDevAttrMultiProcessorCount, dev);
  const int sm_count = sm_counts[slot];
  size_t& smem_opted = smem_opts[slot];
  if (smem_bytes > smem_opted || slot != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem_bytes));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(hmc_kernel, %zu B): %s", smem_bytes, cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
    smem_opted = smem_bytes;
  }
  int bps = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, kThreads, smem_bytes);
  if (e != cudaSuccess || bps < 1) {
    set_error("hmc_kernel does not fit an SM (%zu B shared): %s", smem_bytes, cudaGetErrorString(e));
    return L2HMC_ERR_CUDA;
  }
  p.num_tiles = (p.N + G::CPB - 1) / G::CPB;
  p.one = 1.0f;
  if (p.num_tiles == 0) return L2HMC_OK;
  const int grid = min(p.num_tiles, bps * sm_count);
  kern<<<grid, kThreads, smem_bytes, stream>>>(p);
  return check_launch("hmc_kernel");
}

template <bool MC>
int dispatch(const HmcParams& p, int T, int X, cudaStream_t stream) {
#define L2HMC_SHAPE(TT, XX) \
  if (T == TT && X == XX) return launch<TT, XX, MC>(p, stream);
  L2HMC_SHAPE(4, 4)
  L2HMC_SHAPE(8, 8)
  L2HMC_SHAPE(16, 16)
  L2HMC_SHAPE(32, 32)
  L2HMC_SHAPE(8, 16)
  L2HMC_SHAPE(16, 32)
#undef L2HMC_SHAPE
  set_error("no fused HMC kernel for a %dx%d lattice", T, X);
  return L2HMC_ERR_UNSUPPORTED;
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_u1_hmc_step(const float* x, const float* v, float beta, float eps, const int8_t* dir, int N,
                                 int T, int X, float* x_out, float* v_out, void* stream) {
  if (N == 0) return L2HMC_OK;  // empty batch: nothing to enqueue (pointers may be null)
  L2HMC_REQUIRE(x && v && x_out && v_out, "l2hmc_u1_hmc_step: null pointer");
  L2HMC_REQUIRE(N >= 0 && T > 0 && X > 0, "l2hmc_u1_hmc_step: bad sizes N=%d T=%d X=%d", N, T, X);
  L2HMC_REQUIRE(aligned16(x) && aligned16(v) && aligned16(x_out) && aligned16(v_out),
                "l2hmc_u1_hmc_step: pointers must be 16-byte aligned");
  HmcParams p{};
  p.x = x, p.v_f = v, p.dir = dir, p.beta = beta, p.eps = eps, p.num_steps = 1, p.N = N;
  p.x_post = x_out, p.v_post = v_out;
  return dispatch<false>(p, T, X, static_cast<cudaStream_t>(stream));
}

extern "C" int l2hmc_u1_hmc_mc_step(const l2hmc_hmc_mc_args* a, void* stream) {
  L2HMC_REQUIRE(a, "l2hmc_u1_hmc_mc_step: null args");
  if (a->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(a->x, "l2hmc_u1_hmc_mc_step: null x");
  L2HMC_REQUIRE(a->N >= 0 && a->T > 0 && a->X > 0 && a->num_steps >= 1, "l2hmc_u1_hmc_mc_step: bad sizes");
  L2HMC_REQUIRE(a->v_f || !a->v_b, "l2hmc_u1_hmc_mc_step: v_b given without v_f");
  L2HMC_REQUIRE(aligned16(a->x) && aligned16(a->v_f) && aligned16(a->v_b) && aligned16(a->x_post) &&
                    aligned16(a->v_post) && aligned16(a->x_out),
                "l2hmc_u1_hmc_mc_step: pointers must be 16-byte aligned");
  HmcParams p{};
  p.x = a->x, p.v_f = a->v_f, p.v_b = a->v_b, p.u_dir = a->u_dir, p.u_acc = a->u_acc;
  p.beta = a->beta, p.eps = a->eps, p.num_steps = a->num_steps, p.default_dir = a->default_dir;
  p.wrap = a->wrap_2pi, p.N = a->N, p.seed = a->seed, p.mc_step = a->mc_step, p.chain0 = a->chain0;
  p.mc_step_dev = reinterpret_cast<const unsigned long long*>(a->mc_step_dev);
  p.x_post = a->x_post, p.v_post = a->v_post, p.accept_prob = a->accept_prob, p.x_out = a->x_out;
  p.accepted = a->accepted, p.h_old = a->h_old, p.h_new = a->h_new, p.flag = a->nonfinite_flag;
  p.obs_action = a->obs_action, p.obs_plaq = a->obs_plaq, p.obs_charge = a->obs_charge;
  return dispatch<true>(p, a->T, a->X, static_cast<cudaStream_t>(stream));
}
