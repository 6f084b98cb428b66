// K3 (FFMA path): dense trunk + heads of the S/T/Q networks, with the K2 leapfrog update fused
// into the head GEMM's epilogue.
//
// Replaces (reference l2hmc/...):
//   network/generic_net.py:129-146   GenericNet.call
//   network/conv_net.py:262-280      dense trunk + heads of ConvNet3D.call
//   dynamics/gauge_dynamics.py:486-590  the four sub-updates (epilogue modes MOMENTUM / POSITION)
//
// One tiled fp32 GEMM kernel, four uses:
//   layer 1   h1 = relu([in0 | m*in1] . [Wv ; Wx] + tau.Wt + bv + bx + bt)   two K segments, the
//             position net's `mask * x` input folded into the A-tile load (no mask_mul pass)
//   layer 2   h2 = relu(h1 . Wh + bh)
//   heads     one GEMM over the three head matrices: a CTA owns 32 outputs d of ALL THREE heads, so
//             every thread holds S, T, Q of the same (chain, d) and applies tanh/exp and -- in the
//             update modes -- the leapfrog sub-update and its log-det without S,T,Q touching HBM.
//   grad      d act = d out . W^T of the backward pass (both operands K-contiguous, up to three K
//             segments; ReLU gate, column mask, accumulate) -- l2hmc_sgemm routes those here.
// CTA tile 64 rows x 64 (96 for the heads) columns x 32 K, 256 threads, 4 x 4 (4 x 6) register
// tile, two K slabs of operands in flight in registers, two shared-memory buffers.  K is cut into
// fixed chunks (a function of the K extents only): fmaf in K order inside a chunk, chunk sums added
// in chunk order -- by the CTAs of one thread-block cluster over distributed shared memory when the
// launch has fewer tiles than SMs (small batches are latency-bound: config 2, 128-256 rows), by one
// CTA walking the chunks otherwise; a row's result does not depend on the batch it is launched with.
// This path serves the shapes too small or ragged for the tcgen05 kernel in gemm_tc.cu.
#include <cooperative_groups.h>

#include <climits>
#include <cstdlib>
#include <string>

#include "common.cuh"
#include "nets.cuh"
#include "update_math.cuh"

namespace l2hmc {
namespace {

constexpr int BM = 64, BK = 32, kThreads = 256;
constexpr int AV = BK / 16;    // float4 A loads per thread and K slab
constexpr int kMaxSeg = 3;
constexpr int kMaxChunks = 8;  // K chunks per output tile = portable cluster size

enum Mode { kBiasRelu = 0, kHeadsStq = 1, kHeadsMomentum = 2, kHeadsPosition = 3, kGrad = 4 };

struct Seg {
  const float* A;      // [rows, K] row-major, leading dimension lda
  int lda;
  const float* mask;   // nullable [K]
  int invert;
  const float* mask_alt;   // rows with row_set != 0
  int invert_alt;
  const float* W[3];   // B_KC = false: [K, cols] row-major, leading dimension ldb (three for the heads)
                       // B_KC = true:  [cols, K] row-major (the weight matrix of the forward layer; C = A . W^T)
  int ldb;
  int K;
};

struct DenseParams {
  Seg seg[kMaxSeg];
  int nseg;
  int rows, cols;       // cols = H (layers) or D (heads: per head)
  // K chunking: the flattened (segment, 32-wide slab) sequence is cut into `nchunks` chunks of `sp` slabs; every chunk
  // is summed in K order from zero and the chunk sums are added in chunk order.  `sp` depends on the K extents only, so
  // an output element gets the same bits whatever the batch size or tile position -- whether one CTA walks the chunks
  // (CL = false) or the `nchunks` CTAs of a cluster take one each and add them over distributed shared memory.
  int sp, nchunks;
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
  // kGrad: C[r*ldc + c] (=|+=) scale[c] * (gate[r*ldg + c] > 0) * sum
  float* C;
  long long ldc, ldg;
  int accumulate;
  const float* gate;
  const float* cscale;
  const float* cscale_alt;
  int cinv, cinv_alt;
};

// NG column groups of CW contiguous columns per thread: layers / gradients 1 x 4 (BN = 64), heads 3 x 2 (BN = 96).
// CL: the K chunks of a tile are the CTAs of one cluster (blockIdx.z = chunk).
template <int NG, int CW, int MODE, bool B_KC, bool CL>
__global__ void __launch_bounds__(kThreads) dense_kernel(const DenseParams p) {
  constexpr int GS = 16 * CW;        // columns per group in the tile
  constexpr int BN = NG * GS;
  constexpr int BNP = BN + 4;        // padded row of the B tile / of the accumulator tile
  constexpr int NC = NG * CW;        // accumulator columns per thread
  constexpr int BV = (BK * BN / 4 + kThreads - 1) / kThreads;   // float4 B loads per thread
  constexpr int kAs = BK * (BM + 4), kBs = BK * BNP;
  __shared__ __align__(16) float smem[2 * kAs + 2 * kBs];
  static_assert(BM * BNP <= 2 * kAs + 2 * kBs, "the accumulator tile re-uses the operand buffers");
  float* const As0 = smem;
  float* const As1 = smem + kAs;
  float* const Bs0 = smem + 2 * kAs;
  float* const Bs1 = Bs0 + kBs;
  float* const red = smem;           // [BM][BNP] after the main loop

  // programmatic dependent launch: do not touch memory before the kernels ahead of this one in the stream have
  // completed (a no-op unless launched with the attribute)
  pdl_wait();
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM;
  const int n0 = blockIdx.x * GS;    // first column (per group) of this tile
  const int z = CL ? blockIdx.z : 0;

  // A: AV float4 along K per thread (k offsets ak + 16 q)
  const int ar = tid >> 2, ak = (tid & 3) * 4;
  const bool a_alt = p.row_set != nullptr && m0 + ar < p.rows && p.row_set[m0 + ar] != 0;

  auto load_tile = [&](float (&a_reg)[AV][4], float (&b_reg)[BV][4], const Seg& s, int k0) {
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
        if (B_KC) {       // float4 along K of output column n
          const int n = idx / (BK / 4), k4 = (idx % (BK / 4)) * 4;
          const int gk = k0 + k4, gc = n0 + n;
          if (gc < p.cols) {
            const float* src = s.W[0] + static_cast<size_t>(gc) * s.ldb + gk;
            if (gk + 3 < s.K && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0)) {
              const float4 t = *reinterpret_cast<const float4*>(src);
              b_reg[q][0] = t.x, b_reg[q][1] = t.y, b_reg[q][2] = t.z, b_reg[q][3] = t.w;
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                if (gk + i < s.K) b_reg[q][i] = src[i];
            }
          }
        } else {
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
    }
  };
  auto store_tile = [&](const float (&a_reg)[AV][4], const float (&b_reg)[BV][4], float* As, float* Bs) {
#pragma unroll
    for (int q = 0; q < AV; ++q)
#pragma unroll
      for (int i = 0; i < 4; ++i) As[(ak + 16 * q + i) * (BM + 4) + ar] = a_reg[q][i];
#pragma unroll
    for (int q = 0; q < BV; ++q) {
      const int idx = tid + q * kThreads;
      if (idx < BK * BN / 4) {
        if (B_KC) {
          const int n = idx / (BK / 4), k4 = (idx % (BK / 4)) * 4;
#pragma unroll
          for (int i = 0; i < 4; ++i) Bs[(k4 + i) * BNP + n] = b_reg[q][i];
        } else {
          const int k = idx / (BN / 4), c4 = (idx % (BN / 4)) * 4;
          *reinterpret_cast<float4*>(&Bs[k * BNP + c4]) = make_float4(b_reg[q][0], b_reg[q][1], b_reg[q][2], b_reg[q][3]);
        }
      }
    }
  };

  float acc[4][NC];
  auto multiply = [&](const float* As, const float* Bs) {
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k * (BM + 4) + 4 * ty]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      float b[NC];
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        if (CW == 4) {
          const float4 t = *reinterpret_cast<const float4*>(&Bs[k * BNP + g * GS + 4 * tx]);
          b[g * CW + 0] = t.x, b[g * CW + 1] = t.y, b[g * CW + 2] = t.z, b[g * CW + 3] = t.w;
        } else if (CW == 2) {
          const float2 t = *reinterpret_cast<const float2*>(&Bs[k * BNP + g * GS + 2 * tx]);
          b[g * CW + 0] = t.x, b[g * CW + 1] = t.y;
        } else {
          b[g * CW] = Bs[k * BNP + g * GS + tx];
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < NC; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  };

  // flattened (segment, k0) slab sequence
  int nsl[kMaxSeg], total_slabs = 0;
#pragma unroll
  for (int s = 0; s < kMaxSeg; ++s) {
    nsl[s] = s < p.nseg ? (p.seg[s].K + BK - 1) / BK : 0;
    total_slabs += nsl[s];
  }
  auto load_slab = [&](float (&a_reg)[AV][4], float (&b_reg)[BV][4], int slab) {
    int sg = 0;
#pragma unroll
    for (int s = 0; s < kMaxSeg - 1; ++s)
      if (sg == s && slab >= nsl[s]) slab -= nsl[s], ++sg;
    load_tile(a_reg, b_reg, p.seg[sg], slab * BK);
  };

  float tot[4][NC];
  float ra0[AV][4], rb0[BV][4], ra1[AV][4], rb1[BV][4];
  for (int c = CL ? z : 0; c < (CL ? z + 1 : p.nchunks); ++c) {
    const int s0 = c * p.sp, n = min(total_slabs, s0 + p.sp) - s0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < NC; ++j) acc[i][j] = 0.f;
    // two slabs of operands in flight in registers, two shared-memory buffers: one barrier per slab
    if (n > 0) load_slab(ra0, rb0, s0);
    if (n > 1) load_slab(ra1, rb1, s0 + 1);
    for (int s = 0; s < n; s += 2) {
      store_tile(ra0, rb0, As0, Bs0);
      __syncthreads();
      if (s + 2 < n) load_slab(ra0, rb0, s0 + s + 2);
      multiply(As0, Bs0);
      if (s + 1 < n) {
        store_tile(ra1, rb1, As1, Bs1);
        __syncthreads();
        if (s + 3 < n) load_slab(ra1, rb1, s0 + s + 3);
        multiply(As1, Bs1);
      }
    }
    __syncthreads();                 // every thread is done with the operand buffers
    if (!CL) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < NC; ++j) tot[i][j] = c == 0 ? acc[i][j] : __fadd_rn(tot[i][j], acc[i][j]);
    }
  }

  // the products are done: let the next kernel of the stream get its CTAs resident under this one's reduce / epilogue
  // (it waits at its own griddepcontrol.wait until this grid has completed)
  pdl_trigger();
  // ------------------------------------------------------------------ accumulators -> shared memory
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
      for (int j = 0; j < CW; ++j)
        red[(4 * ty + i) * BNP + g * GS + CW * tx + j] = CL ? acc[i][g * CW + j] : tot[i][g * CW + j];
  namespace cg = cooperative_groups;
  int r_begin = 0, r_end = BM, nsrc = 1;
  const float* src[kMaxChunks];
  src[0] = red;
  if (CL) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
    nsrc = p.nchunks;
    const int rs = (BM + nsrc - 1) / nsrc;       // rows of the tile this rank finishes
    r_begin = min(BM, z * rs), r_end = min(BM, r_begin + rs);
#pragma unroll
    for (int q = 0; q < kMaxChunks; ++q) src[q] = q < nsrc ? cl.map_shared_rank(red, q) : red;
  } else {
    __syncthreads();
  }

  // ------------------------------------------------------------------ epilogue: thread (tx, ty) finishes rows
  // r_begin + ty + 16 t, columns CW tx + j of every group; chunk sums added in chunk order
  const int c0 = n0 + CW * tx;       // first owned column (per group)
  float add[2][CW], bs[CW], bt[CW], bq[CW], es[CW], eq[CW], um[CW], um_alt[CW];
  if (MODE == kBiasRelu) {
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
  } else if (MODE != kGrad) {
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
  }
  const float heps = __fmul_rn(0.5f, p.eps);
  bool bad = false;
  for (int base = r_begin; base < r_end; base += 16) {     // uniform trip count: the log-det shuffles need whole warps
    const int rl = base + ty;
    const int r = m0 + rl;
    const bool rok = rl < r_end && r < p.rows;
    float val[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) val[j] = 0.f;
    if (rok) {
      float part[kMaxChunks][NC];
#pragma unroll
      for (int q = 0; q < kMaxChunks; ++q) {
        if (q < nsrc) {
#pragma unroll
          for (int g = 0; g < NG; ++g) {
            const float* at = src[q] + rl * BNP + g * GS + CW * tx;
            if (CW == 4) {
              const float4 t = *reinterpret_cast<const float4*>(at);
              part[q][g * CW] = t.x, part[q][g * CW + 1] = t.y, part[q][g * CW + 2] = t.z, part[q][g * CW + 3] = t.w;
            } else if (CW == 2) {
              const float2 t = *reinterpret_cast<const float2*>(at);
              part[q][g * CW] = t.x, part[q][g * CW + 1] = t.y;
            } else {
              part[q][g * CW] = *at;
            }
          }
        }
      }
#pragma unroll
      for (int j = 0; j < NC; ++j) val[j] = part[0][j];
#pragma unroll
      for (int q = 1; q < kMaxChunks; ++q)
        if (q < nsrc) {
#pragma unroll
          for (int j = 0; j < NC; ++j) val[j] = __fadd_rn(val[j], part[q][j]);
        }
    }
    const bool alt = p.row_set != nullptr && rok && p.row_set[r] != 0;
    if (MODE == kBiasRelu) {
      if (!rok) continue;
      float o[CW];
#pragma unroll
      for (int j = 0; j < CW; ++j) o[j] = fmaxf(__fadd_rn(val[j], alt ? add[1][j] : add[0][j]), 0.f);
      float* dst = p.out[0] + static_cast<size_t>(r) * p.cols + c0;
      if (CW == 4 && c0 + 3 < p.cols && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
        *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
      } else {
#pragma unroll
        for (int j = 0; j < CW; ++j)
          if (c0 + j < p.cols) dst[j] = o[j];
      }
    } else if (MODE == kGrad) {
      if (!rok) continue;
      const float* cs = alt ? p.cscale_alt : p.cscale;
      const int cinv = alt ? p.cinv_alt : p.cinv;
#pragma unroll
      for (int j = 0; j < CW; ++j) {
        const int c = c0 + j;
        if (c >= p.cols) continue;
        float v = val[j];
        if (p.gate != nullptr && !(p.gate[static_cast<long long>(r) * p.ldg + c] > 0.f)) v = 0.f;
        if (cs != nullptr) {
          const float sc = cs[c];
          v = __fmul_rn(v, cinv ? __fsub_rn(1.f, sc) : sc);
        }
        float* dst = p.C + static_cast<long long>(r) * p.ldc + c;
        if (p.accumulate) v = __fadd_rn(*dst, v);
        *dst = v;
      }
    } else {
      // heads: val[0..CW) = S, [CW..2CW) = T, [2CW..3CW) = Q pre-activations of outputs d = c0 + j
      const int dir = alt ? p.dir_alt : p.dir;
      float ld = 0.f;
#pragma unroll
      for (int j = 0; j < CW; ++j) {
        const int d = c0 + j;
        if (!(rok && d < p.cols)) continue;
        const float S = __fmul_rn(tanhf(__fadd_rn(val[j], bs[j])), es[j]);
        const float Tr = __fadd_rn(val[CW + j], bt[j]);
        float Q = __fadd_rn(val[2 * CW + j], bq[j]);
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
  }
  if (MODE == kHeadsMomentum || MODE == kHeadsPosition) flag_nonfinite(p.flag, bad);
  if (CL) cg::this_cluster().sync();       // this rank's tile stays alive until every rank has read it
}

// logdet[n] += sum_j part[j][n], j in fixed order (deterministic)
__global__ void logdet_reduce_kernel(const float* __restrict__ part, int tiles, int N, float* logdet) {
  pdl_wait();
  pdl_trigger();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int j = 0; j < tiles; ++j) s = __fadd_rn(s, part[static_cast<size_t>(j) * N + n]);
  logdet[n] = __fadd_rn(logdet[n], s);
}

// The K chunking rule: from the K extents only (never the batch size), so that a row's result does not depend on how
// many rows were launched with it.  At least 2 slabs (64 products) per chunk, at most kMaxChunks chunks.
void set_chunks(DenseParams& p) {
  int total = 0;
  for (int s = 0; s < p.nseg; ++s) total += (p.seg[s].K + BK - 1) / BK;
  p.sp = std::max(2, (total + kMaxChunks - 1) / kMaxChunks);
  p.nchunks = std::max(1, (total + p.sp - 1) / p.sp);
}

constexpr int kClusterBelowTiles = 148;   // fewer output tiles than SMs: spread the K chunks of a tile over a cluster

template <int NG, int CW, int MODE, bool B_KC>
int launch_dense(DenseParams p, cudaStream_t st) {
  constexpr int GS = 16 * CW;
  set_chunks(p);
  const int ct = (p.cols + GS - 1) / GS, rt = (p.rows + BM - 1) / BM;
  const bool cl = p.nchunks > 1 && ct * rt < kClusterBelowTiles;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(ct, rt, cl ? p.nchunks : 1), cfg.blockDim = dim3(kThreads), cfg.dynamicSmemBytes = 0, cfg.stream = st;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 1, attr.val.clusterDim.y = 1, attr.val.clusterDim.z = cl ? p.nchunks : 1;
  cudaLaunchAttribute attrs[2];
  attrs[0] = attr;
  attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs, cfg.numAttrs = pdl_enabled() ? 2 : 1;
  const cudaError_t e = cl ? cudaLaunchKernelEx(&cfg, dense_kernel<NG, CW, MODE, B_KC, true>, p)
                           : cudaLaunchKernelEx(&cfg, dense_kernel<NG, CW, MODE, B_KC, false>, p);
  if (e != cudaSuccess) return check_launch("dense_kernel", 0);
  return check_launch("dense_kernel");
}

int launch_layer(const DenseParams& p, cudaStream_t st) { return launch_dense<1, 4, kBiasRelu, false>(p, st); }

}  // namespace

// The per-row log-det partials of the head kernel: one per 32-output column tile.
int heads_col_tiles(int D, int /*N*/) { return (D + 31) / 32; }

// C = A . B^T products of the backward pass with both operands K-contiguous (d act = d out . W^T: l2hmc_gemm_seg with
// a_cs == 1 and b_rs == 1), up to three K segments, through the same chunked kernel as the forward layers.
bool dense_grad_nt_ok(const l2hmc_gemm_args* g) {
  if (g->nseg < 1 || g->nseg > kMaxSeg) return false;
  for (int s = 0; s < g->nseg; ++s) {
    const l2hmc_gemm_seg& sg = g->seg[s];
    if (sg.a_cs != 1 || sg.b_rs != 1 || sg.a_rs > INT_MAX || sg.b_cs > INT_MAX || sg.a_rs < 0 || sg.b_cs < 0) return false;
  }
  return true;
}

int dense_grad_nt(const l2hmc_gemm_args* g, cudaStream_t st) {
  DenseParams p{};
  p.nseg = g->nseg;
  for (int s = 0; s < g->nseg; ++s) {
    const l2hmc_gemm_seg& sg = g->seg[s];
    p.seg[s] = Seg{sg.A, static_cast<int>(sg.a_rs), nullptr, 0, nullptr, 0, {sg.B, nullptr, nullptr},
                   static_cast<int>(sg.b_cs), sg.K};
  }
  p.rows = g->M, p.cols = g->N;
  p.C = g->C, p.ldc = g->ldc, p.ldg = g->ldg, p.accumulate = g->accumulate, p.gate = g->relu_gate;
  p.row_set = g->row_set;
  p.cscale = g->col_scale, p.cinv = g->col_scale_invert;
  p.cscale_alt = g->row_set ? g->col_scale_alt : g->col_scale, p.cinv_alt = g->row_set ? g->col_scale_invert_alt : g->col_scale_invert;
  return launch_dense<1, 4, kGrad, true>(p, st);
}

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
  if (c->mode == L2HMC_NET_STQ) rc = launch_dense<3, 2, kHeadsStq, false>(p, st);
  else if (c->mode == L2HMC_NET_MOMENTUM) rc = launch_dense<3, 2, kHeadsMomentum, false>(p, st);
  else rc = launch_dense<3, 2, kHeadsPosition, false>(p, st);
  if (rc) return rc;
  if (c->mode != L2HMC_NET_STQ && c->logdet) {
    launch_pdl(logdet_reduce_kernel, dim3((N + 255) / 256), dim3(256), 0, st, ld_part, heads_col_tiles(D, N), N, c->logdet);
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
