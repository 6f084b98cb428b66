// One host call = one training step's loss AND gradients (forward with tape + explicit backward sweep).
//
// Replaces the graph the reference builds in gauge_model.py:728-832 (_calc_loss + tf.gradients over
// dynamics/gauge_dynamics.py:195-313, 412-609) -- the same launches, in the same order, as the Python-driven walk of
// l2hmc_b200/training.py (TrainEngine.forward / backward), issued back to back from C++: a Python loop pays ~30 us
// of interpreter + ctypes time per launch, which at config 2 (128 chains, ~290 launches) is the whole step.
// Results are bit-identical to the Python-driven composition (tests/test_gpu_training.py).
//
// Batch layout: the x-chain and the auxiliary z-chain are stacked into R = 2N rows; forward- and backward-going
// chains share every launch (row_set).  Tape per network (2M calls): inputs, h1, h2, S, T, Q, then the activation
// gradients; weight gradients are deferred split-K GEMMs over the whole tape (see train.cu).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <initializer_list>
#include <vector>

#include "common.cuh"
#include "nets.cuh"

namespace l2hmc {
namespace {

__global__ void __launch_bounds__(256) train_row_set_kernel(const float* __restrict__ u_dir, int R, uint8_t* row_set) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < R) row_set[n] = u_dir[n] > 0.5f ? 0 : 1;
}

// v0[n,:] = row_set[n] ? v_b[n,:] : v_f[n,:]   (the reference's exact 0/1 blend of its two momentum draws)
__global__ void __launch_bounds__(128) pick_rows_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                        const uint8_t* __restrict__ row_set, int D, float* out) {
  const int row = blockIdx.x;
  const float* src = (row_set[row] ? b : a) + static_cast<size_t>(row) * D;
  for (int d = threadIdx.x; d < D; d += 128) out[static_cast<size_t>(row) * D + d] = src[d];
}

// (cos, sin, 1) per chain and network call: the t_layer input of the call and the "input" of every bias.  Call j of a
// network belongs to leapfrog step j / 2; forward chains read time row `step`, backward chains row M - 1 - step.  The
// (cos, sin) rows are a HOST array (l2hmc_train_args.ts): they travel as a kernel parameter.
constexpr int kTauTableSteps = 128;
struct TauTable {
  float v[2 * kTauTableSteps];
  int M;        // direct == 0: v holds all M rows, call j = j0 + blockIdx.y
  int direct;   // direct == 1: v[0..1] = the forward chains' row, v[2..3] = the backward chains' row of call j0
  int j0;
};
__global__ void __launch_bounds__(256) tau_rows_kernel(const uint8_t* __restrict__ row_set, const TauTable tb, int R,
                                                       float* out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= R) return;
  const int j = tb.j0 + blockIdx.y, step = j >> 1;
  const bool alt = row_set[n] != 0;
  const int row = tb.direct ? (alt ? 1 : 0) : (alt ? tb.M - 1 - step : step);
  float* o = out + (static_cast<size_t>(j) * R + n) * 3;
  o[0] = tb.v[2 * row], o[1] = tb.v[2 * row + 1], o[2] = 1.f;
}

struct Carve {
  char* p;
  size_t used;
  template <typename T>
  T* take(size_t count) {
    const size_t bytes = (count * sizeof(T) + 255) & ~static_cast<size_t>(255);
    T* r = p ? reinterpret_cast<T*>(p + used) : nullptr;
    used += bytes;
    return r;
  }
};

struct NetTape {
  int K, H, D;
  float *in0, *in1, *h1, *h2, *S, *T, *Q, *dA, *dz2, *dz1, *ps, *pq, *tau1;
};

struct Plan {
  int N, R, D, M, calls;
  bool nets, conv;
  float *X, *V;          // states: [(2M+1)][R,D]
  float* G;              // forces of the momentum calls [2M][R,D] (generic nets: lives in the VNet tape)
  uint8_t* row_set;
  float *sld, *h_old, *h_new, *p, *a, *g_p, *deps, *gx_post, *force, *gx, *gv, *dg, *scratch;
  NetTape xt, vt;
  void* net_ws;
  size_t net_ws_bytes;
  void *xpack, *vpack;   // TS32 copies of the dense weights (tcgen05 forward), re-packed every step
  size_t xpack_bytes, vpack_bytes;
  bool tc_forward;
  float *fg0, *fg1;
  float* part[4];        // conv partial rows: XNet.x, XNet.v, VNet.x, VNet.v
  int prow, pcol;
  // conv routing tapes (l2hmc_conv3d_tape): per network and input stream, one slot per call
  float* cp1[2][2];      // [net: 0 = XNet, 1 = VNet][stream: 0 = in0, 1 = in1][2M calls][R][tape_elems]
  uint8_t* ca1[2][2];
  uint8_t* ca2[2][2];    // [..][R][K]
  size_t tape_elems;
  void* gemm_ws;
  size_t gemm_ws_bytes;
  // chained tcgen05 backward products (tc_gemm_chain): d z2 and d z1 of the current call in the tiled TF32 form
  bool chain;
  float *dz2_ts, *dz1_ts;
  size_t total;
};

int ksplit_for(int M, int N, int K) {
  const int tiles = ((M + 63) / 64) * ((N + 63) / 64);
  return std::max(1, std::min(64, std::min((592 + tiles - 1) / tiles, K / 64)));   // ~4 CTAs per SM (measured: 16 tiles, K = 2560: 19.7 us at 19 ranges, 17.1 at 32-40)
}

void carve_tape(Carve& c, NetTape& t, const l2hmc_net_weights* w, size_t Rt) {
  t.K = w->K, t.H = w->H, t.D = w->D;
  t.in0 = c.take<float>(Rt * t.K), t.in1 = c.take<float>(Rt * t.K);
  t.h1 = c.take<float>(Rt * t.H), t.h2 = c.take<float>(Rt * t.H);
  t.S = c.take<float>(Rt * t.D), t.T = c.take<float>(Rt * t.D), t.Q = c.take<float>(Rt * t.D);
  t.dA = c.take<float>(Rt * 3 * t.D);
  t.dz2 = c.take<float>(Rt * t.H), t.dz1 = c.take<float>(Rt * t.H);
  t.ps = c.take<float>(Rt * t.D), t.pq = c.take<float>(Rt * t.D);
  t.tau1 = c.take<float>(Rt * 3);
}

Plan make_plan(const l2hmc_train_args* a, void* base) {
  Plan pl{};
  pl.N = a->N, pl.R = 2 * a->N, pl.D = 2 * a->T * a->X, pl.M = a->M, pl.calls = 4 * a->M;
  pl.nets = a->xnet != nullptr;
  pl.conv = a->xconv_in0 != nullptr;
  const size_t R = pl.R, D = pl.D, M = pl.M, RD = R * D;
  Carve c{static_cast<char*>(base), 0};
  pl.X = c.take<float>((2 * M + 1) * RD);
  pl.V = c.take<float>((2 * M + 1) * RD);
  pl.G = (pl.nets && !pl.conv) ? nullptr : c.take<float>(2 * M * RD);
  pl.row_set = c.take<uint8_t>(R);
  pl.sld = c.take<float>(R), pl.h_old = c.take<float>(R), pl.h_new = c.take<float>(R), pl.p = c.take<float>(R);
  pl.a = c.take<float>(R), pl.g_p = c.take<float>(R), pl.deps = c.take<float>(R);
  pl.gx_post = c.take<float>(RD), pl.force = c.take<float>(RD), pl.gx = c.take<float>(RD), pl.gv = c.take<float>(RD);
  pl.dg = c.take<float>(RD);
  pl.scratch = c.take<float>(1024);
  if (pl.nets) {
    const size_t Rt = 2 * M * R;
    carve_tape(c, pl.xt, a->xnet, Rt);
    carve_tape(c, pl.vt, a->vnet, Rt);
    pl.net_ws_bytes = std::max(l2hmc_net_workspace_bytes(a->xnet->K, a->xnet->H, a->xnet->D, pl.R),
                               l2hmc_net_workspace_bytes(a->vnet->K, a->vnet->H, a->vnet->D, pl.R));
    pl.net_ws = c.take<char>(pl.net_ws_bytes);
    // forward layers on tcgen05 when the stacked batch fills MMA tiles (same rule as l2hmc_net_run's AUTO path, from
    // 1024 rows: below that the FFMA layers are as fast and the step is launch-bound anyway)
    auto fits = [](const l2hmc_net_weights* w) { return w->K >= 64 && w->H >= 128 && w->D >= 64 && w->D % 16 == 0; };
    static const int tc_min_rows = getenv("L2HMC_TC_MIN_ROWS") ? atoi(getenv("L2HMC_TC_MIN_ROWS")) : 1024;   // diagnostics
    pl.tc_forward = pl.R >= tc_min_rows && fits(a->xnet) && fits(a->vnet);
    if (pl.tc_forward) {
      pl.xpack_bytes = l2hmc_net_pack_bytes(a->xnet->K, a->xnet->H, a->xnet->D);
      pl.vpack_bytes = l2hmc_net_pack_bytes(a->vnet->K, a->vnet->H, a->vnet->D);
      pl.xpack = c.take<char>(pl.xpack_bytes), pl.vpack = c.take<char>(pl.vpack_bytes);
    }
    size_t gw = 0;
    for (const l2hmc_net_weights* w : {a->xnet, a->vnet}) {
      const int K = static_cast<int>(Rt);
      auto need = [&](int Mm, int Nn, int nseg, int Kk, int ksplit) {     // what l2hmc_gemm will ask for this shape
        l2hmc_gemm_args g{};
        g.M = Mm, g.N = Nn, g.nseg = nseg;
        for (int s = 0; s < nseg; ++s) g.seg[s].K = Kk;
        gw = std::max(gw, l2hmc_gemm_workspace_bytes(&g, ksplit));
      };
      auto deferred = [&](int Mm, int Nn) { need(Mm, Nn, 1, K, ksplit_for(Mm, Nn, K)); };
      deferred(w->H, w->D), deferred(1, w->D), deferred(w->H, w->H), deferred(1, w->H), deferred(w->K, w->H);
      deferred(2, w->H);
      need(pl.R, w->H, 3, w->D, 0), need(pl.R, w->H, 1, w->H, 0), need(pl.R, w->K, 1, w->H, 0);   // per-call GEMMs
      gw = std::max(gw, col_sum_multi_workspace_bytes(K, std::max(w->H, w->D), 7));
    }
    pl.gemm_ws_bytes = gw;
    pl.gemm_ws = c.take<char>(gw);
    // the per-call backward products d z2 = dA . Wheads^T, d z1 = d z2 . Wh^T, d in = d z1 . Win^T run on tcgen05 for
    // both networks: each hands its result to the next already tiled (no operand-conversion pass over d z2 / d z1)
    static const bool chain_on = !(getenv("L2HMC_TC_CHAIN") && getenv("L2HMC_TC_CHAIN")[0] == '0');   // diagnostics
    pl.chain = chain_on;
    size_t ts = 0;
    for (const l2hmc_net_weights* w : {a->xnet, a->vnet}) {
      auto tc = [&](int Mm, int Nn, int nseg, int Kk) {
        l2hmc_gemm_args g{};
        g.M = Mm, g.N = Nn, g.nseg = nseg;
        for (int s = 0; s < nseg; ++s) g.seg[s].K = Kk;
        return tc_gemm_chainable(&g);
      };
      pl.chain = pl.chain && tc(pl.R, w->H, 3, w->D) && tc(pl.R, w->H, 1, w->H) && tc(pl.R, w->K, 1, w->H);
      ts = std::max(ts, tc_gemm_ts_floats(pl.R, w->H));
    }
    if (pl.chain) pl.dz2_ts = c.take<float>(ts), pl.dz1_ts = c.take<float>(ts);
    if (pl.conv) {
      pl.fg0 = c.take<float>(R * a->xnet->K), pl.fg1 = c.take<float>(R * a->xnet->K);
      pl.prow = l2hmc_conv3d_bwd_rows(pl.R), pl.pcol = l2hmc_conv3d_bwd_cols(a->xconv_in0->F);
      for (int i = 0; i < 4; ++i) pl.part[i] = c.take<float>(static_cast<size_t>(pl.prow) * pl.pcol);
      pl.tape_elems = static_cast<size_t>(l2hmc_conv3d_tape_elems(a->T, a->X, a->xconv_in0->F));
      const size_t calls = 2 * M;
      for (int net = 0; net < 2; ++net)
        for (int sidx = 0; sidx < 2; ++sidx) {
          pl.cp1[net][sidx] = c.take<float>(calls * R * pl.tape_elems);
          pl.ca1[net][sidx] = c.take<uint8_t>(calls * R * pl.tape_elems);
          pl.ca2[net][sidx] = c.take<uint8_t>(calls * R * static_cast<size_t>(a->xnet->K));
        }
    }
  }
  pl.total = c.used + 256;
  return pl;
}

#define L2HMC_TRY(expr)            \
  do {                             \
    const int rc_ = (expr);        \
    if (rc_ != L2HMC_OK) return rc_; \
  } while (0)

struct Ctx {
  const l2hmc_train_args* a;
  Plan pl;
  cudaStream_t st;
  void* stream;
};

int gemm(const Ctx& c, int M, int N, std::initializer_list<l2hmc_gemm_seg> segs, float* C, long long ldc, int accumulate,
         const float* gate, long long ldg, const float* cs, int inv, const float* cs_alt, int inv_alt, bool sets, int ksplit,
         const float* a_ts = nullptr, float* out_ts = nullptr) {
  l2hmc_gemm_args g{};
  g.M = M, g.N = N, g.nseg = static_cast<int>(segs.size());
  int i = 0;
  for (const auto& s : segs) g.seg[i++] = s;
  g.C = C, g.ldc = ldc, g.accumulate = accumulate, g.relu_gate = gate, g.ldg = ldg;
  g.col_scale = cs, g.col_scale_invert = inv;
  g.row_set = sets ? c.pl.row_set : nullptr, g.col_scale_alt = cs_alt, g.col_scale_invert_alt = inv_alt;
  if (a_ts != nullptr || out_ts != nullptr)       // a link of a tcgen05 chain (Plan::chain)
    return tc_gemm_chain(&g, a_ts, out_ts, c.pl.gemm_ws, c.pl.gemm_ws_bytes, c.st);
  return l2hmc_gemm(&g, ksplit, c.pl.gemm_ws, c.pl.gemm_ws_bytes, c.stream);
}

int plain_gemm(const Ctx& c, int M, int N, l2hmc_gemm_seg s, float* C, long long ldc, int accumulate = 0, int ksplit = 0,
               const float* a_ts = nullptr) {
  return gemm(c, M, N, {s}, C, ldc, accumulate, nullptr, 0, nullptr, 0, nullptr, 0, false, ksplit, a_ts);
}

int col_sum(const Ctx& c, const float* B, long long ld, int rows, int cols, float* out) {
  return l2hmc_col_sum(B, ld, rows, cols, out, c.pl.gemm_ws, c.pl.gemm_ws_bytes, c.stream);
}

int net_forward(const Ctx& c, const l2hmc_net_weights* w, const void* packed, NetTape& t, int j, const float* tau,
                const float* tau_alt) {
  const Plan& pl = c.pl;
  const size_t R = pl.R, off = static_cast<size_t>(j) * R;
  l2hmc_net_weights ww = *w;
  ww.packed = packed;
  l2hmc_net_call n{};
  n.mode = L2HMC_NET_STQ;
  n.in0 = t.in0 + off * t.K, n.in1 = t.in1 + off * t.K;
  n.tau_cos = tau[0], n.tau_sin = tau[1], n.tau_cos_alt = tau_alt[0], n.tau_sin_alt = tau_alt[1];
  n.row_set = pl.row_set, n.dir = n.dir_alt = 1, n.N = pl.R;
  n.S = t.S + off * t.D, n.T = t.T + off * t.D, n.Q = t.Q + off * t.D;
  n.h1_out = t.h1 + off * t.H, n.h2_out = t.h2 + off * t.H;
  n.workspace = pl.net_ws, n.workspace_bytes = pl.net_ws_bytes;
  n.path = packed ? L2HMC_NET_PATH_TCGEN05 : L2HMC_NET_PATH_FFMA;
  return l2hmc_net_run(&ww, &n, c.stream);
}

// dA[j] -> dz2[j] -> dz1[j]
int net_backward(const Ctx& c, const l2hmc_net_weights* w, NetTape& t, int j) {
  const size_t R = c.pl.R, off = static_cast<size_t>(j) * R;
  const int H = t.H, D = t.D;
  float* dA = t.dA + off * 3 * D;
  L2HMC_TRY(gemm(c, c.pl.R, H,
                 {l2hmc_gemm_seg{dA, 3LL * D, 1, w->w_s, 1, D, D}, l2hmc_gemm_seg{dA + D, 3LL * D, 1, w->w_tr, 1, D, D},
                  l2hmc_gemm_seg{dA + 2 * D, 3LL * D, 1, w->w_q, 1, D, D}},
                 t.dz2 + off * H, H, 0, t.h2 + off * H, H, nullptr, 0, nullptr, 0, false, 0, nullptr,
                 c.pl.chain ? c.pl.dz2_ts : nullptr));
  return gemm(c, c.pl.R, H, {l2hmc_gemm_seg{t.dz2 + off * H, H, 1, w->w_h, 1, H, H}}, t.dz1 + off * H, H, 0,
              t.h1 + off * H, H, nullptr, 0, nullptr, 0, false, 0, c.pl.chain ? c.pl.dz2_ts : nullptr,
              c.pl.chain ? c.pl.dz1_ts : nullptr);
}

int weight_grads(const Ctx& c, const NetTape& t, const l2hmc_net_grads* g) {
  const int Rt = 2 * c.pl.M * c.pl.R, H = t.H, D = t.D, K = t.K;
  auto ks = [&](int M, int N) { return ksplit_for(M, N, Rt); };
  float* wk[3] = {g->w_s, g->w_tr, g->w_q};
  for (int s = 0; s < 3; ++s)
    L2HMC_TRY(plain_gemm(c, H, D, l2hmc_gemm_seg{t.h2, 1, H, t.dA + s * D, 3LL * D, 1, Rt}, wk[s], D, 0, ks(H, D)));
  L2HMC_TRY(plain_gemm(c, H, H, l2hmc_gemm_seg{t.h1, 1, H, t.dz2, H, 1, Rt}, g->w_h, H, 0, ks(H, H)));
  L2HMC_TRY(plain_gemm(c, K, H, l2hmc_gemm_seg{t.in0, 1, K, t.dz1, H, 1, Rt}, g->w_in0, H, 0, ks(K, H)));
  L2HMC_TRY(plain_gemm(c, K, H, l2hmc_gemm_seg{t.in1, 1, K, t.dz1, H, 1, Rt}, g->w_in1, H, 0, ks(K, H)));
  L2HMC_TRY(plain_gemm(c, 2, H, l2hmc_gemm_seg{t.tau1, 1, 3, t.dz1, H, 1, Rt}, g->w_t, H, 0, ks(2, H)));
  // the seven bias / coefficient gradients (batch sums of activation gradients) in two launches; the three first-layer
  // biases add into the same sum
  const ColSumJob jobs[7] = {
      {t.dA, 3LL * D, D, {g->b_s, nullptr, nullptr}},        {t.dA + D, 3LL * D, D, {g->b_tr, nullptr, nullptr}},
      {t.dA + 2 * D, 3LL * D, D, {g->b_q, nullptr, nullptr}}, {t.dz2, H, H, {g->b_h, nullptr, nullptr}},
      {t.dz1, H, H, {g->b_in0, g->b_in1, g->b_t}},            {t.ps, D, D, {g->coeff_s, nullptr, nullptr}},
      {t.pq, D, D, {g->coeff_q, nullptr, nullptr}}};
  return col_sum_multi(jobs, 7, Rt, c.pl.gemm_ws, c.pl.gemm_ws_bytes, c.st);
}

int conv_weight_grads(const Ctx& c, const NetTape& t, const float* part, const l2hmc_conv3d_grads* g, int F) {
  (void)t;
  const int rows = c.pl.prow, cols = c.pl.pcol;
  float* dst[4] = {g->w1, g->b1, g->w2, g->b2};
  const int n[4] = {18 * F, F, 16 * F * F, 2 * F};
  int off = 0;
  for (int i = 0; i < 4; ++i) {
    L2HMC_TRY(col_sum(c, part + off, cols, rows, n[i], dst[i]));
    off += n[i];
  }
  return L2HMC_OK;
}

// tape slot of (net, input stream, call j)
l2hmc_conv3d_tape conv_tape(const Ctx& c, int net, int sidx, int j) {
  const Plan& pl = c.pl;
  const size_t off = static_cast<size_t>(j) * pl.R;
  return l2hmc_conv3d_tape{pl.cp1[net][sidx] + off * pl.tape_elems, pl.ca1[net][sidx] + off * pl.tape_elems,
                           pl.ca2[net][sidx] + off * static_cast<size_t>(c.a->xnet->K)};
}

int conv_fwd(const Ctx& c, const l2hmc_conv3d_weights* w, const float* raw, float* feat, const l2hmc_conv3d_tape& tape,
             const float* mask, int inv, const float* mask_alt, int inv_alt) {
  return l2hmc_conv3d_features_tape(w, raw, mask, inv, mask_alt, inv_alt, mask ? c.pl.row_set : nullptr, c.pl.R, c.a->T,
                                    c.a->X, feat, &tape, c.stream);
}

int conv_bwd(const Ctx& c, const l2hmc_conv3d_weights* w, const float* raw, const float* g_feat, float* g_in, float* part,
             const float* mask, int inv, const float* mask_alt, int inv_alt, const l2hmc_conv3d_tape& tape) {
  l2hmc_conv3d_bwd_args b{};
  b.tape = &tape;
  b.w = w, b.in = raw, b.N = c.pl.R, b.T = c.a->T, b.X = c.a->X;
  b.mask = mask, b.mask_invert = inv, b.mask_alt = mask_alt, b.mask_invert_alt = inv_alt;
  b.row_set = mask ? c.pl.row_set : nullptr;
  b.g_feat = g_feat, b.g_in = g_in, b.partial = part;
  return l2hmc_conv3d_backward(&b, c.stream);
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

static int validate_train(const l2hmc_train_args* a) {
  L2HMC_REQUIRE(a, "l2hmc_u1_train_grads: null arguments");
  L2HMC_REQUIRE(a->N > 0 && a->T > 0 && a->X > 0 && a->M > 0, "l2hmc_u1_train_grads: bad sizes N=%d T=%d X=%d M=%d", a->N,
                a->T, a->X, a->M);
  L2HMC_REQUIRE((a->xnet == nullptr) == (a->vnet == nullptr), "l2hmc_u1_train_grads: xnet and vnet go together");
  const bool conv = a->xconv_in0 != nullptr;
  L2HMC_REQUIRE(conv == (a->xconv_in1 != nullptr) && conv == (a->vconv_in0 != nullptr) && conv == (a->vconv_in1 != nullptr),
                "l2hmc_u1_train_grads: the four conv weight sets must be given together");
  L2HMC_REQUIRE(!conv || a->xnet, "l2hmc_u1_train_grads: conv weights without networks");
  return L2HMC_OK;
}

extern "C" size_t l2hmc_u1_train_workspace_bytes(const l2hmc_train_args* a) {
  if (validate_train(a) != L2HMC_OK) return 0;
  return make_plan(a, nullptr).total;
}

extern "C" int l2hmc_u1_train_grads(const l2hmc_train_args* a, void* stream) {
  L2HMC_TRY(validate_train(a));
  L2HMC_REQUIRE(a->masks && a->ts && a->x && a->z && a->v_f && a->v_b && a->u_dir && a->u_acc,
                "l2hmc_u1_train_grads: null input");
  L2HMC_REQUIRE(a->loss_out && a->per_chain && a->x_out && a->px && a->g_eps, "l2hmc_u1_train_grads: null output");
  const bool nets = a->xnet != nullptr, conv = a->xconv_in0 != nullptr;
  L2HMC_REQUIRE(!nets || (a->xnet_g && a->vnet_g), "l2hmc_u1_train_grads: gradient pointers of the networks missing");
  L2HMC_REQUIRE(!conv || (a->xconv_in0_g && a->xconv_in1_g && a->vconv_in0_g && a->vconv_in1_g),
                "l2hmc_u1_train_grads: gradient pointers of the conv weights missing");
  const int N = a->N, T = a->T, X = a->X, M = a->M, D = 2 * T * X, R = 2 * N;
  L2HMC_REQUIRE(!nets || (a->xnet->D == D && a->vnet->D == D), "l2hmc_u1_train_grads: network output width != 2TX");
  L2HMC_REQUIRE(a->workspace && a->workspace_bytes >= l2hmc_u1_train_workspace_bytes(a),
                "l2hmc_u1_train_grads: workspace too small (%zu < %zu bytes)", a->workspace_bytes,
                l2hmc_u1_train_workspace_bytes(a));
  L2HMC_REQUIRE((reinterpret_cast<uintptr_t>(a->workspace) & 255u) == 0, "l2hmc_u1_train_grads: workspace must be "
                "256-byte aligned");
  Ctx c{a, make_plan(a, a->workspace), static_cast<cudaStream_t>(stream), stream};
  Plan& pl = c.pl;
  cudaStream_t st = c.st;
  const size_t RD = static_cast<size_t>(R) * D, ND = static_cast<size_t>(N) * D;
  const float eps = a->eps, beta = a->beta;
  auto Xs = [&](int k) { return pl.X + static_cast<size_t>(k) * RD; };
  auto Vs = [&](int k) { return pl.V + static_cast<size_t>(k) * RD; };
  auto copy = [&](float* dst, const float* src, size_t n) {
    return cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st) == cudaSuccess ? L2HMC_OK : L2HMC_ERR_CUDA;
  };

  // ------------------------------------------------------------------ forward with tape
  L2HMC_TRY(copy(Xs(0), a->x, ND));
  L2HMC_TRY(copy(Xs(0) + ND, a->z, ND));
  train_row_set_kernel<<<(R + 255) / 256, 256, 0, st>>>(a->u_dir, R, pl.row_set);
  pick_rows_kernel<<<R, 128, 0, st>>>(a->v_f, a->v_b, pl.row_set, D, Vs(0));
  L2HMC_TRY(check_launch("pick_rows_kernel", 2));
  if (nets) {       // the time-embedding rows of all 2M calls of each network
    for (NetTape* t : {&pl.xt, &pl.vt}) {
      TauTable tb{};
      if (M <= kTauTableSteps) {
        for (int i = 0; i < 2 * M; ++i) tb.v[i] = a->ts[i];
        tb.M = M, tb.direct = 0, tb.j0 = 0;
        tau_rows_kernel<<<dim3((R + 255) / 256, 2 * M), 256, 0, st>>>(pl.row_set, tb, R, t->tau1);
        L2HMC_TRY(check_launch("tau_rows_kernel"));
      } else {
        for (int j = 0; j < 2 * M; ++j) {
          const int step = j >> 1, ib = M - 1 - step;
          tb.v[0] = a->ts[2 * step], tb.v[1] = a->ts[2 * step + 1], tb.v[2] = a->ts[2 * ib], tb.v[3] = a->ts[2 * ib + 1];
          tb.M = M, tb.direct = 1, tb.j0 = j;
          tau_rows_kernel<<<dim3((R + 255) / 256, 1), 256, 0, st>>>(pl.row_set, tb, R, t->tau1);
          L2HMC_TRY(check_launch("tau_rows_kernel"));
        }
      }
    }
  }
  cudaMemsetAsync(pl.sld, 0, R * sizeof(float), st);
  if (nets && pl.tc_forward) {      // the weights moved since the last step: refresh their TS32 copies
    L2HMC_TRY(l2hmc_net_pack(a->xnet, pl.xpack, pl.xpack_bytes, stream));
    L2HMC_TRY(l2hmc_net_pack(a->vnet, pl.vpack, pl.vpack_bytes, stream));
  }
  struct Rec { int kind, j, xi, vi, inv, inv_alt, step, ib; float* g; };
  std::vector<Rec> recs;
  recs.reserve(4 * M);
  int jx = 0, jv = 0, xi = 0, vi = 0;
  // generic nets: the network inputs are copies / mask products of the states.  Every update kernel writes them for
  // the calls that consume its result (UpdExtra); only the start state needs launches of its own.
  const bool dense_in = nets && !conv;
  auto slot = [&](float* base, int j, int K) { return base + static_cast<size_t>(j) * R * K; };
  if (dense_in) {
    L2HMC_TRY(copy(pl.vt.in0, Xs(0), RD));
    L2HMC_TRY(l2hmc_mask_mul_sets(Xs(0), a->masks, 0, a->masks + static_cast<size_t>(M - 1) * D, 1, pl.row_set, R, D,
                                  pl.xt.in1, stream));
  }
  for (int step = 0; step < M; ++step) {
    const int ib = M - step - 1;
    const float* tau = a->ts + 2 * step;
    const float* tau_alt = a->ts + 2 * ib;
    const float* row_f = a->masks + static_cast<size_t>(step) * D;
    const float* row_b = a->masks + static_cast<size_t>(ib) * D;
    const int kinds[4] = {0, 1, 1, 0}, invs[4] = {0, 0, 1, 0}, invs_alt[4] = {0, 1, 0, 0};
    for (int q = 0; q < 4; ++q) {
      if (kinds[q] == 0) {          // momentum update
        float* g = (nets && !conv) ? pl.vt.in1 + static_cast<size_t>(jv) * R * pl.vt.K : pl.G + static_cast<size_t>(jv) * RD;
        recs.push_back(Rec{0, jv, xi, vi, 0, 0, step, ib, g});
        L2HMC_TRY(l2hmc_u1_action_force(Xs(xi), beta, R, T, X, nullptr, g, stream));
        const float *S = nullptr, *Tr = nullptr, *Q = nullptr;
        if (nets) {
          NetTape& t = pl.vt;
          const size_t off = static_cast<size_t>(jv) * R;
          if (conv) {   // momentum_fn([x, grad, t]): x through conv_v*, the force through conv_x*
            L2HMC_TRY(conv_fwd(c, a->vconv_in0, Xs(xi), t.in0 + off * t.K, conv_tape(c, 1, 0, jv), nullptr, 0, nullptr, 0));
            L2HMC_TRY(conv_fwd(c, a->vconv_in1, g, t.in1 + off * t.K, conv_tape(c, 1, 1, jv), nullptr, 0, nullptr, 0));
          }             // (generic nets: in0 = x was written by the update that produced x)
          L2HMC_TRY(net_forward(c, a->vnet, pl.tc_forward ? pl.vpack : nullptr, t, jv, tau, tau_alt));
          S = t.S + off * D, Tr = t.T + off * D, Q = t.Q + off * D;
        }
        UpdExtra ex{};
        if (dense_in && q == 0)       // the new v is position_fn's first input in both position updates of this step
          ex.copy[0] = slot(pl.xt.in0, jx, pl.xt.K), ex.copy[1] = slot(pl.xt.in0, jx + 1, pl.xt.K);
        L2HMC_TRY(train_update_fwd(0, nullptr, Vs(vi), g, S, Tr, Q, nullptr, 0, nullptr, 0, pl.row_set, eps, a->eps_dev, 1,
                                   -1, R, D, Vs(vi + 1), pl.sld, a->nonfinite_flag, &ex, stream));
        ++vi, ++jv;
      } else {                      // position update
        recs.push_back(Rec{1, jx, xi, vi, invs[q], invs_alt[q], step, ib, nullptr});
        const float *S = nullptr, *Tr = nullptr, *Q = nullptr;
        if (nets) {
          NetTape& t = pl.xt;
          const size_t off = static_cast<size_t>(jx) * R;
          if (conv) {   // position_fn([v, mask * x, t]): v through conv_v*, m*x through conv_x*
            L2HMC_TRY(conv_fwd(c, a->xconv_in0, Vs(vi), t.in0 + off * t.K, conv_tape(c, 0, 0, jx), nullptr, 0, nullptr, 0));
            L2HMC_TRY(conv_fwd(c, a->xconv_in1, Xs(xi), t.in1 + off * t.K, conv_tape(c, 0, 1, jx), row_f, invs[q], row_b,
                               invs_alt[q]));
          }             // (generic nets: in0 = v and in1 = m * x were written by the updates that produced v and x)
          L2HMC_TRY(net_forward(c, a->xnet, pl.tc_forward ? pl.xpack : nullptr, t, jx, tau, tau_alt));
          S = t.S + off * D, Tr = t.T + off * D, Q = t.Q + off * D;
        }
        UpdExtra ex{};
        if (dense_in && q == 1) {     // the new x, masked for the second position update of this step
          ex.masked = slot(pl.xt.in1, jx + 1, pl.xt.K);
          ex.nmask = row_f, ex.ninv = invs[2], ex.nmask_alt = row_b, ex.ninv_alt = invs_alt[2];
        } else if (dense_in) {        // q == 2: the new x is momentum_fn's first input in the next two momentum updates
          ex.copy[0] = slot(pl.vt.in0, jv, pl.vt.K);       // ... and, masked, position_fn's second input in the next step
          if (step + 1 < M) {
            ex.copy[1] = slot(pl.vt.in0, jv + 1, pl.vt.K);
            ex.masked = slot(pl.xt.in1, jx + 1, pl.xt.K);
            ex.nmask = a->masks + static_cast<size_t>(step + 1) * D, ex.ninv = invs[1];
            ex.nmask_alt = a->masks + static_cast<size_t>(M - step - 2) * D, ex.ninv_alt = invs_alt[1];
          }
        }
        L2HMC_TRY(train_update_fwd(1, Xs(xi), Vs(vi), nullptr, S, Tr, Q, row_f, invs[q], row_b, invs_alt[q], pl.row_set, eps,
                                   a->eps_dev, 1, -1, R, D, Xs(xi + 1), pl.sld, a->nonfinite_flag, &ex, stream));
        ++xi, ++jx;
      }
    }
  }
  const float *xM = Xs(xi), *vM = Vs(vi);
  L2HMC_TRY(l2hmc_u1_hamiltonian(Xs(0), Vs(0), beta, R, T, X, pl.h_old, stream));
  L2HMC_TRY(l2hmc_u1_hamiltonian(xM, vM, beta, R, T, X, pl.h_new, stream));
  L2HMC_TRY(l2hmc_accept_prob(pl.h_old, pl.h_new, pl.sld, R, pl.p, a->nonfinite_flag, stream));
  L2HMC_TRY(l2hmc_mh_select(a->x, xM, nullptr, nullptr, nullptr, pl.p, nullptr, nullptr, a->u_acc, N, D, 0, nullptr, nullptr,
                            nullptr, a->x_out, a->accepted, stream));
  L2HMC_TRY(l2hmc_u1_loss(&a->loss, a->x, xM, a->z, a->x_out, pl.p, pl.p + N, N, T, X, a->per_chain, a->x_dq, a->loss_out,
                          stream));
  L2HMC_TRY(copy(a->px, pl.p, N));
  if (a->pz) L2HMC_TRY(copy(a->pz, pl.p + N, N));
  if (a->x_prop) L2HMC_TRY(copy(a->x_prop, xM, ND));
  if (a->forward_only) return L2HMC_OK;

  // ------------------------------------------------------------------ backward sweep
  cudaMemsetAsync(pl.gx_post, 0, RD * sizeof(float), st);      // z rows: z_prop is not in the loss
  cudaMemsetAsync(pl.deps, 0, R * sizeof(float), st);
  if (conv)
    for (int i = 0; i < 4; ++i) cudaMemsetAsync(pl.part[i], 0, static_cast<size_t>(pl.prow) * pl.pcol * sizeof(float), st);
  L2HMC_TRY(l2hmc_u1_loss_bwd(&a->loss, a->x, xM, a->z, pl.p, pl.p + N, N, T, X, pl.gx_post, pl.g_p, pl.g_p + N, stream));
  L2HMC_TRY(l2hmc_accept_prob_bwd(pl.h_old, pl.h_new, pl.sld, pl.g_p, R, pl.a, stream));
  L2HMC_TRY(l2hmc_u1_action_force(xM, beta, R, T, X, nullptr, pl.force, stream));
  L2HMC_TRY(l2hmc_traj_bwd_init(pl.a, pl.force, vM, pl.gx_post, R, D, pl.gx, pl.gv, stream));
  for (int r = static_cast<int>(recs.size()) - 1; r >= 0; --r) {
    const Rec& rec = recs[r];
    const float* row_f = a->masks + static_cast<size_t>(rec.step) * D;
    const float* row_b = a->masks + static_cast<size_t>(rec.ib) * D;
    l2hmc_update_bwd_args u{};
    u.N = R, u.D = D, u.eps = eps, u.eps_dev = a->eps_dev, u.dir = 1, u.dir_alt = -1, u.row_set = pl.row_set;
    u.v = Vs(rec.vi), u.gv = pl.gv, u.gld = pl.a, u.deps = pl.deps;
    const l2hmc_net_weights* w = nullptr;
    NetTape* t = nullptr;
    const size_t off = static_cast<size_t>(rec.j) * R;
    if (nets) {
      w = rec.kind == 0 ? a->vnet : a->xnet;
      t = rec.kind == 0 ? &pl.vt : &pl.xt;
      u.S = t->S + off * D, u.T = t->T + off * D, u.Q = t->Q + off * D;
      u.coeff_s = w->coeff_s, u.coeff_q = w->coeff_q;
      u.dA = t->dA + off * 3 * D, u.ps = t->ps + off * D, u.pq = t->pq + off * D;
    }
    if (rec.kind == 0) {
      u.position = 0, u.g = rec.g, u.dg = pl.dg;
      L2HMC_TRY(l2hmc_update_bwd(&u, stream));
      if (nets) {
        L2HMC_TRY(net_backward(c, w, *t, rec.j));
        const int H = t->H, K = t->K;
        const float* dz1 = t->dz1 + off * H;
        const float* dz1_ts = pl.chain ? pl.dz1_ts : nullptr;
        if (conv) {
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in0, 1, H, H}, pl.fg0, K, 0, 0, dz1_ts));
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in1, 1, H, H}, pl.fg1, K, 0, 0, dz1_ts));
          L2HMC_TRY(conv_bwd(c, a->vconv_in0, Xs(rec.xi), pl.fg0, pl.gx, pl.part[3], nullptr, 0, nullptr, 0,
                             conv_tape(c, 1, 0, rec.j)));
          L2HMC_TRY(conv_bwd(c, a->vconv_in1, rec.g, pl.fg1, pl.dg, pl.part[2], nullptr, 0, nullptr, 0,
                             conv_tape(c, 1, 1, rec.j)));
        } else {
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in0, 1, H, H}, pl.gx, D, 1, 0, dz1_ts));
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in1, 1, H, H}, pl.dg, D, 1, 0, dz1_ts));
        }
      }
      L2HMC_TRY(l2hmc_u1_force_hvp(Xs(rec.xi), pl.dg, beta, R, T, X, pl.gx, stream));
    } else {
      u.position = 1, u.x = Xs(rec.xi), u.gx = pl.gx;
      u.mask = row_f, u.invert = rec.inv, u.mask_alt = row_b, u.invert_alt = rec.inv_alt;
      L2HMC_TRY(l2hmc_update_bwd(&u, stream));
      if (nets) {
        L2HMC_TRY(net_backward(c, w, *t, rec.j));
        const int H = t->H, K = t->K;
        const float* dz1 = t->dz1 + off * H;
        const float* dz1_ts = pl.chain ? pl.dz1_ts : nullptr;
        if (conv) {
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in0, 1, H, H}, pl.fg0, K, 0, 0, dz1_ts));
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in1, 1, H, H}, pl.fg1, K, 0, 0, dz1_ts));
          L2HMC_TRY(conv_bwd(c, a->xconv_in0, Vs(rec.vi), pl.fg0, pl.gv, pl.part[1], nullptr, 0, nullptr, 0,
                             conv_tape(c, 0, 0, rec.j)));
          L2HMC_TRY(conv_bwd(c, a->xconv_in1, Xs(rec.xi), pl.fg1, pl.gx, pl.part[0], row_f, rec.inv, row_b, rec.inv_alt,
                             conv_tape(c, 0, 1, rec.j)));
        } else {
          L2HMC_TRY(plain_gemm(c, R, K, l2hmc_gemm_seg{dz1, H, 1, w->w_in0, 1, H, H}, pl.gv, D, 1, 0, dz1_ts));
          L2HMC_TRY(gemm(c, R, K, {l2hmc_gemm_seg{dz1, H, 1, w->w_in1, 1, H, H}}, pl.gx, D, 1, nullptr, 0, row_f, rec.inv,
                         row_b, rec.inv_alt, true, 0, dz1_ts));
        }
      }
    }
  }
  // ------------------------------------------------------------------ parameter gradients
  L2HMC_TRY(l2hmc_reduce_sum(pl.deps, R, 0, a->g_eps, pl.scratch, stream));
  if (nets) {
    L2HMC_TRY(weight_grads(c, pl.xt, a->xnet_g));
    if (conv) {
      L2HMC_TRY(conv_weight_grads(c, pl.xt, pl.part[0], a->xconv_in1_g, a->xconv_in0->F));
      L2HMC_TRY(conv_weight_grads(c, pl.xt, pl.part[1], a->xconv_in0_g, a->xconv_in0->F));
    }
    if (a->first_bucket_event != nullptr &&
        cudaEventRecord(static_cast<cudaEvent_t>(a->first_bucket_event), st) != cudaSuccess) {
      set_error("l2hmc_u1_train_grads: cudaEventRecord(first_bucket_event): %s", cudaGetErrorString(cudaGetLastError()));
      return L2HMC_ERR_CUDA;
    }
    L2HMC_TRY(weight_grads(c, pl.vt, a->vnet_g));
    if (conv) {
      L2HMC_TRY(conv_weight_grads(c, pl.vt, pl.part[2], a->vconv_in1_g, a->xconv_in0->F));
      L2HMC_TRY(conv_weight_grads(c, pl.vt, pl.part[3], a->vconv_in0_g, a->xconv_in0->F));
    }
  }
  return L2HMC_OK;
}
