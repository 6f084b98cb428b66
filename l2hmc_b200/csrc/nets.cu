// K3 (FFMA path): dense trunk + heads of the S/T/Q networks, with the K2 leapfrog update fused
// into the head GEMM's epilogue.
//
// Replaces (reference l2hmc/...):
//   network/generic_net.py:129-146   GenericNet.call
//   network/conv_net.py:262-280      dense trunk + heads of ConvNet3D.call
//   dynamics/gauge_dynamics.py:486-590  the four sub-updates (epilogue modes MOMENTUM / POSITION)
//
// One tiled fp32 GEMM kernel, three uses:
//   layer 1   h1 = relu([in0 | m*in1] . [Wv ; Wx] + tau.Wt + bv + bx + bt)   two K segments, the
//             position net's `mask * x` input folded into the A-tile load (no mask_mul pass)
//   layer 2   h2 = relu(h1 . Wh + bh)
//   heads     one GEMM over the three head matrices: a CTA owns 32 outputs d of ALL THREE heads, so
//             every thread holds S, T, Q of the same (chain, d) and applies tanh/exp and -- in the
//             update modes -- the leapfrog sub-update and its log-det without S,T,Q touching HBM.
// CTA tile 64 rows x 64 (96 for the heads) columns x 16 K, 256 threads, 4 x 4 (4 x 6) register
// tile, operands prefetched into registers while the previous K slab is multiplied.  Accumulation
// is fmaf in K order (fp32); this path serves the shapes too small or ragged for the tcgen05
// kernel in gemm_tc.cu (config 2: 128 chains x 512 hidden; tail rows).
#include "common.cuh"
#include "nets.cuh"
#include "update_math.cuh"

namespace l2hmc {
namespace {

constexpr int BM = 64, BK = 32, kThreads = 256;
constexpr int AV = BK / 16;   // float4 A loads per thread and K slab

enum Mode { kBiasRelu = 0, kHeadsStq = 1, kHeadsMomentum = 2, kHeadsPosition = 3 };

struct Seg {
  const float* A;      // [rows, K] row-major, leading dimension lda
  int lda;
  const float* mask;   // nullable [K]
  int invert;
  const float* mask_alt;   // rows with row_set != 0
  int invert_alt;
  const float* W[3];   // [K, cols] row-major, leading dimension ldb (three for the heads)
  int ldb;
  int K;
};

struct DenseParams {
  Seg seg[2];
  int nseg;
  int rows, cols;       // cols = H (layers) or D (heads: per head)
  const float* bias[3]; // layers: up to three biases summed; heads: b_s, b_tr, b_q
  const float* wt;      // t_layer kernel [2, cols] (nullable)
  float tau0, tau1;
  const uint8_t* row_set;   // nullable [rows]: 1 -> the *_alt parameters (mixed-direction batches)
  float tau0_alt, tau1_alt;
  const float* umask_alt;
  int uinvert_alt, dir_alt;
  const float* coeff_s; // heads
  const float* coeff_q;
  int tanh_on_q;
  float* out[3];        // layers: out[0] [rows, cols]; kHeadsStq: S, T, Q [rows, cols]
  // update modes
  const float* ux;
  const float* uv;
  const float* ug;
  const float* umask;
  int uinvert;
  float eps;
  int dir;
  float* uout;
  float* ld_part;       // [col tiles][rows]
  int32_t* flag;
};

// NG column groups of CW contiguous columns per thread: layers 1 x 4 (BN = 64), heads 3 x 2 (BN = 96)
template <int NG, int CW, int MODE>
__global__ void __launch_bounds__(kThreads) dense_kernel(const DenseParams p) {
  constexpr int GS = 16 * CW;        // columns per group in the tile
  constexpr int BN = NG * GS;
  constexpr int NC = NG * CW;        // accumulator columns per thread
  constexpr int BV = (BK * BN / 4 + kThreads - 1) / kThreads;   // float4 B loads per thread
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM;
  const int n0 = blockIdx.x * GS;    // first column (per group) of this tile

  float acc[4][NC];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < NC; ++j) acc[i][j] = 0.f;

  // A: AV float4 along K per thread (k offsets ak + 16 q)
  const int ar = tid >> 2, ak = (tid & 3) * 4;
  const bool a_alt = p.row_set != nullptr && m0 + ar < p.rows && p.row_set[m0 + ar] != 0;
  float a_reg[AV][4];
  float b_reg[BV][4];

  auto load_tile = [&](const Seg& s, int k0) {
    const int gr = m0 + ar;
#pragma unroll
    for (int q = 0; q < AV; ++q) {
#pragma unroll
      for (int i = 0; i < 4; ++i) a_reg[q][i] = 0.f;
      if (gr < p.rows) {
        const int kq = k0 + ak + 16 * q;
        const float* src = s.A + static_cast<size_t>(gr) * s.lda + kq;
        if (kq + 3 < s.K && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0)) {
          const float4 t = *reinterpret_cast<const float4*>(src);
          a_reg[q][0] = t.x, a_reg[q][1] = t.y, a_reg[q][2] = t.z, a_reg[q][3] = t.w;
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (kq + i < s.K) a_reg[q][i] = src[i];
        }
        const float* mask = a_alt ? s.mask_alt : s.mask;
        if (mask != nullptr) {
          const int invert = a_alt ? s.invert_alt : s.invert;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (kq + i < s.K) {
              const float m = mask[kq + i];
              a_reg[q][i] = __fmul_rn(invert ? __fsub_rn(1.f, m) : m, a_reg[q][i]);
            }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < BV; ++q) {
      const int idx = tid + q * kThreads;       // float4 index inside the B tile
#pragma unroll
      for (int i = 0; i < 4; ++i) b_reg[q][i] = 0.f;
      if (idx < BK * BN / 4) {
        const int k = idx / (BN / 4), c4 = (idx % (BN / 4)) * 4;
        const int g = c4 / GS, c = c4 % GS;
        const int gk = k0 + k, gc = n0 + c;
        if (gk < s.K) {
          const float* src = s.W[g] + static_cast<size_t>(gk) * s.ldb + gc;
          if (gc + 3 < p.cols && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0)) {
            const float4 t = *reinterpret_cast<const float4*>(src);
            b_reg[q][0] = t.x, b_reg[q][1] = t.y, b_reg[q][2] = t.z, b_reg[q][3] = t.w;
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (gc + i < p.cols) b_reg[q][i] = src[i];
          }
        }
      }
    }
  };
  auto store_tile = [&]() {
#pragma unroll
    for (int q = 0; q < AV; ++q)
#pragma unroll
      for (int i = 0; i < 4; ++i) As[ak + 16 * q + i][ar] = a_reg[q][i];
#pragma unroll
    for (int q = 0; q < BV; ++q) {
      const int idx = tid + q * kThreads;
      if (idx < BK * BN / 4) {
        const int k = idx / (BN / 4), c4 = (idx % (BN / 4)) * 4;
        *reinterpret_cast<float4*>(&Bs[k][c4]) = make_float4(b_reg[q][0], b_reg[q][1], b_reg[q][2], b_reg[q][3]);
      }
    }
  };

  // flattened (segment, k0) slab sequence
  int total_slabs = 0;
  for (int s = 0; s < p.nseg; ++s) total_slabs += (p.seg[s].K + BK - 1) / BK;
  int cur_seg = 0, cur_k0 = 0;
  auto advance = [&]() {
    cur_k0 += BK;
    if (cur_k0 >= p.seg[cur_seg].K) ++cur_seg, cur_k0 = 0;
  };
  if (total_slabs > 0) load_tile(p.seg[0], 0);
  for (int slab = 0; slab < total_slabs; ++slab) {
    store_tile();
    __syncthreads();
    advance();
    if (slab + 1 < total_slabs) load_tile(p.seg[cur_seg], cur_k0);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][4 * ty]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      float b[NC];
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        if (CW == 4) {
          const float4 t = *reinterpret_cast<const float4*>(&Bs[k][g * GS + 4 * tx]);
          b[g * CW + 0] = t.x, b[g * CW + 1] = t.y, b[g * CW + 2] = t.z, b[g * CW + 3] = t.w;
        } else if (CW == 2) {
          const float2 t = *reinterpret_cast<const float2*>(&Bs[k][g * GS + 2 * tx]);
          b[g * CW + 0] = t.x, b[g * CW + 1] = t.y;
        } else {
          b[g * CW] = Bs[k][g * GS + tx];
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < NC; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  // ------------------------------------------------------------------ epilogue
  const int c0 = n0 + CW * tx;       // first owned column (per group)
  if (MODE == kBiasRelu) {
    float add[2][CW];
#pragma unroll
    for (int j = 0; j < CW; ++j) {
      const int c = c0 + j;
      float s = 0.f;
      if (c < p.cols) {
        s = p.bias[0][c];
        if (p.bias[1]) s = __fadd_rn(s, p.bias[1][c]);
      }
      add[0][j] = add[1][j] = s;
      if (c < p.cols && p.wt) {
        const float w0 = p.wt[c], w1 = p.wt[p.cols + c], b2 = p.bias[2][c];
        add[0][j] = __fadd_rn(s, __fadd_rn(__fadd_rn(__fmul_rn(p.tau0, w0), __fmul_rn(p.tau1, w1)), b2));
        add[1][j] = __fadd_rn(s, __fadd_rn(__fadd_rn(__fmul_rn(p.tau0_alt, w0), __fmul_rn(p.tau1_alt, w1)), b2));
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = m0 + 4 * ty + i;
      if (r >= p.rows) continue;
      const int set = (p.row_set != nullptr && p.row_set[r] != 0) ? 1 : 0;
      float o[CW];
#pragma unroll
      for (int j = 0; j < CW; ++j) o[j] = fmaxf(__fadd_rn(acc[i][j], set ? add[1][j] : add[0][j]), 0.f);
      float* dst = p.out[0] + static_cast<size_t>(r) * p.cols + c0;
      if (CW == 4 && c0 + 3 < p.cols && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
        *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
      } else {
#pragma unroll
        for (int j = 0; j < CW; ++j)
          if (c0 + j < p.cols) dst[j] = o[j];
      }
    }
  } else {
    // heads: acc[i][0..1] = S, [2..3] = T, [4..5] = Q pre-activations of outputs d = c0, c0 + 1
    float bs[CW], bt[CW], bq[CW], es[CW], eq[CW], um[CW], um_alt[CW];
#pragma unroll
    for (int j = 0; j < CW; ++j) {
      const int d = c0 + j;
      const bool ok = d < p.cols;
      bs[j] = ok ? p.bias[0][d] : 0.f;
      bt[j] = ok ? p.bias[1][d] : 0.f;
      bq[j] = ok ? p.bias[2][d] : 0.f;
      es[j] = ok ? expf(p.coeff_s[d]) : 1.f;
      eq[j] = ok ? expf(p.coeff_q[d]) : 1.f;
      um[j] = um_alt[j] = 0.f;
      if (MODE == kHeadsPosition && ok) {
        um[j] = p.uinvert ? __fsub_rn(1.f, p.umask[d]) : p.umask[d];
        if (p.row_set) um_alt[j] = p.uinvert_alt ? __fsub_rn(1.f, p.umask_alt[d]) : p.umask_alt[d];
      }
    }
    const float heps = __fmul_rn(0.5f, p.eps);
    bool bad = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = m0 + 4 * ty + i;
      const bool rok = r < p.rows;
      const bool alt = p.row_set != nullptr && rok && p.row_set[r] != 0;
      const int dir = alt ? p.dir_alt : p.dir;
      float ld = 0.f;
#pragma unroll
      for (int j = 0; j < CW; ++j) {
        const int d = c0 + j;
        if (!(rok && d < p.cols)) continue;
        const float S = __fmul_rn(tanhf(__fadd_rn(acc[i][j], bs[j])), es[j]);
        const float Tr = __fadd_rn(acc[i][CW + j], bt[j]);
        float Q = __fadd_rn(acc[i][2 * CW + j], bq[j]);
        if (p.tanh_on_q) Q = tanhf(Q);
        Q = __fmul_rn(Q, eq[j]);
        const size_t at = static_cast<size_t>(r) * p.cols + d;
        if (MODE == kHeadsStq) {
          p.out[0][at] = S, p.out[1][at] = Tr, p.out[2][at] = Q;
        } else if (MODE == kHeadsMomentum) {
          const float v = p.uv[at], g = p.ug[at];
          p.uout[at] = dir > 0 ? momentum_elem<1>(v, g, S, Tr, Q, p.eps, heps, ld, bad)
                               : momentum_elem<-1>(v, g, S, Tr, Q, p.eps, heps, ld, bad);
        } else {
          const float x = p.ux[at], v = p.uv[at];
          const float m = alt ? um_alt[j] : um[j];
          p.uout[at] = dir > 0 ? position_elem<1>(x, v, S, Tr, Q, m, p.eps, ld, bad)
                               : position_elem<-1>(x, v, S, Tr, Q, m, p.eps, ld, bad);
        }
      }
      if (MODE != kHeadsStq) {
        // the 16 lanes that share this row hold the tile's 32 outputs: fixed shuffle tree
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) ld += __shfl_xor_sync(0xffffffffu, ld, o);
        if (tx == 0 && rok && p.ld_part) p.ld_part[static_cast<size_t>(blockIdx.x) * p.rows + r] = ld;
      }
    }
    if (MODE != kHeadsStq) flag_nonfinite(p.flag, bad);
  }
}

// logdet[n] += sum_j part[j][n], j in fixed order (deterministic)
__global__ void logdet_reduce_kernel(const float* __restrict__ part, int tiles, int N, float* logdet) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int j = 0; j < tiles; ++j) s = __fadd_rn(s, part[static_cast<size_t>(j) * N + n]);
  logdet[n] = __fadd_rn(logdet[n], s);
}

template <int NG, int CW, int MODE>
int launch_dense(const DenseParams& p, cudaStream_t st) {
  constexpr int GS = 16 * CW;
  dim3 grid((p.cols + GS - 1) / GS, (p.rows + BM - 1) / BM);
  dense_kernel<NG, CW, MODE><<<grid, kThreads, 0, st>>>(p);
  return check_launch("dense_kernel");
}

// Layers: 64-column tiles; when those leave most SMs idle (config 2: 256 stacked rows x 512 columns = 32 tiles),
// 32-column tiles.  Every output element sums the same products in the same K order either way (bit-identical).
int launch_layer(const DenseParams& p, cudaStream_t st) {
  const int wide_tiles = ((p.cols + 63) / 64) * ((p.rows + BM - 1) / BM);
  return wide_tiles < 74 ? launch_dense<1, 2, kBiasRelu>(p, st) : launch_dense<1, 4, kBiasRelu>(p, st);
}

}  // namespace

// Heads: a CTA owns 32 outputs of all three heads; when that leaves most SMs idle (config 2: 256 stacked rows x 128
// outputs = 16 CTAs), 16 outputs.  The per-row log-det partials are one per column tile either way.
static bool heads_narrow(int D, int N) { return ((D + 31) / 32) * ((N + BM - 1) / BM) < 74; }
int heads_col_tiles(int D, int N) { return heads_narrow(D, N) ? (D + 15) / 16 : (D + 31) / 32; }

// Layers 1, 2 and the heads on the FFMA path (also called by gemm_tc.cu for shapes it rejects).
int net_run_simt(const l2hmc_net_weights* w, const l2hmc_net_call* c, cudaStream_t st) {
  const int N = c->N, K = w->K, H = w->H, D = w->D;
  float* ws = static_cast<float*>(c->workspace);
  float* h1 = c->h1_out ? c->h1_out : ws;
  float* h2 = c->h2_out ? c->h2_out : ws + static_cast<size_t>(N) * H;
  float* ld_part = ws + 2 * static_cast<size_t>(N) * H;

  DenseParams p{};
  p.nseg = 2;
  p.seg[0] = Seg{c->in0, K, nullptr, 0, nullptr, 0, {w->w_in0, nullptr, nullptr}, H, K};
  p.seg[1] = Seg{c->in1, K, c->in1_mask, c->in1_mask_invert, c->in1_mask_alt, c->in1_mask_invert_alt,
                 {w->w_in1, nullptr, nullptr}, H, K};
  p.rows = N, p.cols = H;
  p.bias[0] = w->b_in0, p.bias[1] = w->b_in1, p.bias[2] = w->b_t;
  p.wt = w->w_t, p.tau0 = c->tau_cos, p.tau1 = c->tau_sin;
  p.row_set = c->row_set, p.tau0_alt = c->tau_cos_alt, p.tau1_alt = c->tau_sin_alt;
  p.out[0] = h1;
  int rc = launch_layer(p, st);
  if (rc) return rc;

  p = DenseParams{};
  p.nseg = 1;
  p.seg[0] = Seg{h1, H, nullptr, 0, nullptr, 0, {w->w_h, nullptr, nullptr}, H, H};
  p.rows = N, p.cols = H;
  p.bias[0] = w->b_h;
  p.out[0] = h2;
  rc = launch_layer(p, st);
  if (rc) return rc;

  p = DenseParams{};
  p.nseg = 1;
  p.seg[0] = Seg{h2, H, nullptr, 0, nullptr, 0, {w->w_s, w->w_tr, w->w_q}, D, H};
  p.rows = N, p.cols = D;
  p.bias[0] = w->b_s, p.bias[1] = w->b_tr, p.bias[2] = w->b_q;
  p.coeff_s = w->coeff_s, p.coeff_q = w->coeff_q, p.tanh_on_q = w->tanh_on_q;
  p.out[0] = c->S, p.out[1] = c->T, p.out[2] = c->Q;
  p.ux = c->x, p.uv = c->v, p.ug = c->g, p.umask = c->mask, p.uinvert = c->mask_invert;
  p.eps = c->eps, p.dir = c->dir, p.uout = c->out, p.flag = c->nonfinite_flag;
  p.row_set = c->row_set, p.umask_alt = c->mask_alt, p.uinvert_alt = c->mask_invert_alt, p.dir_alt = c->dir_alt;
  p.ld_part = c->logdet ? ld_part : nullptr;
  if (heads_narrow(D, N)) {
    if (c->mode == L2HMC_NET_STQ) rc = launch_dense<3, 1, kHeadsStq>(p, st);
    else if (c->mode == L2HMC_NET_MOMENTUM) rc = launch_dense<3, 1, kHeadsMomentum>(p, st);
    else rc = launch_dense<3, 1, kHeadsPosition>(p, st);
  } else {
    if (c->mode == L2HMC_NET_STQ) rc = launch_dense<3, 2, kHeadsStq>(p, st);
    else if (c->mode == L2HMC_NET_MOMENTUM) rc = launch_dense<3, 2, kHeadsMomentum>(p, st);
    else rc = launch_dense<3, 2, kHeadsPosition>(p, st);
  }
  if (rc) return rc;
  if (c->mode != L2HMC_NET_STQ && c->logdet) {
    logdet_reduce_kernel<<<(N + 255) / 256, 256, 0, st>>>(ld_part, heads_col_tiles(D, N), N, c->logdet);
    rc = check_launch("logdet_reduce_kernel");
  }
  return rc;
}

int net_validate(const l2hmc_net_weights* w, const l2hmc_net_call* c) {
  L2HMC_REQUIRE(w && c, "l2hmc_net_run: null argument");
  L2HMC_REQUIRE(w->K > 0 && w->H > 0 && w->D > 0 && c->N >= 0, "l2hmc_net_run: bad sizes K=%d H=%d D=%d N=%d", w->K,
                w->H, w->D, c->N);
  L2HMC_REQUIRE(w->w_in0 && w->b_in0 && w->w_in1 && w->b_in1 && w->w_t && w->b_t && w->w_h && w->b_h && w->w_s &&
                    w->b_s && w->w_tr && w->b_tr && w->w_q && w->b_q && w->coeff_s && w->coeff_q,
                "l2hmc_net_run: null weight pointer");
  if (c->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(c->in0 && c->in1, "l2hmc_net_run: null input");
  L2HMC_REQUIRE(c->workspace && c->workspace_bytes >= l2hmc_net_workspace_bytes(w->K, w->H, w->D, c->N),
                "l2hmc_net_run: workspace too small (%zu < %zu bytes)", c->workspace_bytes,
                l2hmc_net_workspace_bytes(w->K, w->H, w->D, c->N));
  L2HMC_REQUIRE(aligned16(c->workspace), "l2hmc_net_run: workspace must be 16-byte aligned");
  switch (c->mode) {
    case L2HMC_NET_STQ:
      L2HMC_REQUIRE(c->S && c->T && c->Q, "l2hmc_net_run: null S/T/Q output");
      break;
    case L2HMC_NET_MOMENTUM:
      L2HMC_REQUIRE(c->v && c->g && c->out, "l2hmc_net_run: momentum update needs v, g, out");
      L2HMC_REQUIRE(c->dir == 1 || c->dir == -1, "l2hmc_net_run: dir must be +1 or -1");
      break;
    case L2HMC_NET_POSITION:
      L2HMC_REQUIRE(c->x && c->v && c->mask && c->out, "l2hmc_net_run: position update needs x, v, mask, out");
      L2HMC_REQUIRE(c->dir == 1 || c->dir == -1, "l2hmc_net_run: dir must be +1 or -1");
      break;
    default:
      set_error("l2hmc_net_run: unknown mode %d", c->mode);
      return L2HMC_ERR_INVALID;
  }
  L2HMC_REQUIRE((c->h1_out == nullptr) == (c->h2_out == nullptr), "l2hmc_net_run: h1_out and h2_out go together");
  if (c->row_set != nullptr) {
    L2HMC_REQUIRE((c->in1_mask == nullptr) == (c->in1_mask_alt == nullptr),
                  "l2hmc_net_run: in1_mask and in1_mask_alt must be given together");
    if (c->mode != L2HMC_NET_STQ)
      L2HMC_REQUIRE(c->dir_alt == 1 || c->dir_alt == -1, "l2hmc_net_run: dir_alt must be +1 or -1");
    if (c->mode == L2HMC_NET_POSITION) L2HMC_REQUIRE(c->mask_alt, "l2hmc_net_run: row_set given without mask_alt");
  }
  return L2HMC_OK;
}

}  // namespace l2hmc

using namespace l2hmc;

extern "C" size_t l2hmc_net_workspace_bytes(int K, int H, int D, int N) {
  if (K <= 0 || H <= 0 || D <= 0 || N <= 0) return 16;
  // h1, h2 [N,H] (fp32 row-major on the FFMA path; tiled hi/lo pairs on the tcgen05 path need
  // 2x and 128-row padding) + log-det partials [D/32][N]
  const size_t simt = 2 * static_cast<size_t>(N) * H + static_cast<size_t>((D + 15) / 16) * N;   // log-det partials at the finer column tiling (monotonic in N)
  return sizeof(float) * std::max(simt, tc_workspace_floats(K, H, D, N)) + 256;
}

extern "C" int l2hmc_net_run(const l2hmc_net_weights* w, const l2hmc_net_call* c, void* stream) {
  const int rc = net_validate(w, c);
  if (rc != L2HMC_OK || c->N == 0) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (c->path == L2HMC_NET_PATH_FFMA) return net_run_simt(w, c, st);
  const bool tc = c->path == L2HMC_NET_PATH_TCGEN05 ? tc_supported(w, c) : tc_eligible(w, c);
  if (c->path == L2HMC_NET_PATH_TCGEN05 && !tc) {
    set_error("l2hmc_net_run: shape K=%d H=%d D=%d N=%d (packed=%p) has no tcgen05 path", w->K, w->H, w->D, c->N,
              w->packed);
    return L2HMC_ERR_UNSUPPORTED;
  }
  return tc ? net_run_tc(w, c, st) : net_run_simt(w, c, st);
}
