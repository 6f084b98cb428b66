// K3 (tensor-core path): the S/T/Q networks' dense layers on tcgen05 with 3xTF32 split precision.
//
// Replaces the same reference code as nets.cu (network/generic_net.py:129-146,
// network/conv_net.py:262-280, and -- in the head epilogue -- dynamics/gauge_dynamics.py:486-590)
// for the shapes where the layers are dense contractions (chains >= 256, K and H in the hundreds:
// BASELINE configs 3 and 4).
//
// Precision.  The north star asks links / log-dets within 1e-5 of the fp32 reference, which plain
// TF32 (10-bit mantissa) misses by two orders of magnitude.  Every operand is therefore split as
// a = hi + lo (hi = a rounded to TF32, lo = a - hi rounded to TF32) and the product is
// formed from three tensor-core MMAs accumulating in fp32 in TMEM:
//       A.B  ~=  A_hi.B_lo + A_lo.B_hi + A_hi.B_hi          (error ~ 2^-22 relative per term)
//
// Layout ("TS32").  Operands live in HBM already in the shared-memory image the MMA consumes, so a
// tile is ONE contiguous 1-D bulk copy (cp.async.bulk, no tensor map):
//   element (r, k) of a K-major matrix with R rows:
//     float offset = ((kb*RG + rg)*2 + part)*256 + kc*32 + (r%8)*4 + k%4
//     kb = k/32, kc = (k%32)/4, rg = r/8, RG = R/8, part = 0 (hi) / 1 (lo)
//   i.e. 8-row x 16-byte core matrices (128 B, contiguous), K-chunks 128 B apart (LBO), hi and lo
//   of a row group 1 KB apart, row groups 2 KB apart (SBO), one 32-wide K block after the other.
//   UMMA descriptors: K-major, SWIZZLE_NONE.
// Weights are packed once (l2hmc_net_pack); activations are written in this form by the producing
// kernel's epilogue (hidden layers) or by a small converter (network inputs, mask folded in).
//
// Kernel.  Persistent, one CTA per SM, 320 threads, warp-specialised:
//   warps 0-7  epilogue: tcgen05.ld finished K chunks (thread = chain, half the tile's columns) and
//              add them into fp32 running sums in registers; at the end of the tile bias/relu ->
//              hi/lo split -> TS32 store, or, for the heads, tanh/exp + the leapfrog sub-update
//   warp 8     bulk-copy producer (one lane) + TMEM allocation
//   warp 9     MMA issuer (one lane): per 32-wide K block 4 K-steps x 3 products of
//              tcgen05.mma.cta_group::1.kind::tf32, M = 128, N = BN
//   smem ring of 2-3 stages (A 32 KB + B BN*256 B), full/empty mbarriers; two accumulator stages in
//   TMEM (tmem_full / tmem_empty), one per in-flight 128-wide K chunk.
#include <cstdlib>
#include <string>

#include <climits>

#include "common.cuh"
#include "nets.cuh"
#include "tc_ptx.cuh"
#include "update_math.cuh"

namespace l2hmc {
namespace {

constexpr int kEpiWarps = 8, kProducerWarp = 8, kMmaWarp = 9;
constexpr int kTcThreads = (kEpiWarps + 2) * 32;
constexpr int TC_BM = 128;
constexpr uint32_t kATileBytes = TC_BM * 32 * 4 * 2;   // 128 rows x 32 k x (hi, lo)

enum TcMode { kTcLayer = 0, kTcHeadsStq = 1, kTcHeadsMomentum = 2, kTcHeadsPosition = 3, kTcPlain = 4 };

struct TcParams {
  const float* A;      // TS32, a_rg row groups
  const float* B;      // TS32 (weights^T), b_rg row groups
  int a_rg, b_rg, KB;
  int m_tiles, n_tiles;
  int rows;            // valid chains
  int cols;            // valid outputs (H for layers, D per head)
  // layer epilogue
  const float* bias;   // [cols] effective bias
  const float* bias_alt;   // rows with row_set != 0 (first layer of a mixed-direction batch)
  const uint8_t* row_set;  // nullable [rows]
  float* out_ts;       // next layer's A operand (TS32), out_rg row groups, out_kb K blocks
  int out_rg, out_kb;
  float* out_rm;       // nullable: the same activations row-major fp32 [rows, cols] (training tape)
  // heads epilogue
  const float* b_s;
  const float* b_tr;
  const float* b_q;
  const float* coeff_s;
  const float* coeff_q;
  int tanh_on_q;
  float* S;
  float* T;
  float* Q;
  const float* ux;
  const float* uv;
  const float* ug;
  const float* umask;
  int uinvert;
  const float* umask_alt;
  int uinvert_alt;
  float eps;
  int dir, dir_alt;
  float* uout;
  float* ld_part;
  int32_t* flag;
  // plain epilogue (kTcPlain): C[r][n] (=|+=) scale[n] * gate[r][n] * product, row-major fp32
  float* C;
  long long ldc;
  int accumulate;
  const float* gate;
  long long ldg;
  const float* cscale;
  const float* cscale_alt;
  int cinv, cinv_alt;
  // split-K (kTcPlain only): work item = (tile, z); z covers K blocks [z * kb_per, min(KB, (z + 1) * kb_per)) and writes
  // its raw sums to partial[z][rows][cols]; the epilogue runs in the reduce stage (fixed z order).  ksplit <= 1: off.
  int ksplit, kb_per;
  float* partial;
  int debug_noload;    // diagnostics (L2HMC_TC_DEBUG_NOLOAD=1): operand tiles are fetched only while the ring first
                       // fills; later K blocks re-use stale shared memory -> wrong numbers, pure MMA / epilogue timing
};

// tcgen05 / TMEM PTX wrappers: tc_ptx.cuh
// K-major, no swizzle: start address, LBO = 128 B (next 16-byte K chunk), SBO = 2048 B (next 8-row group)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  return static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu) | (static_cast<uint64_t>(128u >> 4) << 16) |
         (static_cast<uint64_t>(2048u >> 4) << 32) | (1ULL << 46);
}

// ------------------------------------------------------------------ the GEMM kernel
// Tensor-core accumulation truncates (round-toward-zero) every time an MMA adds into the TMEM
// accumulator; over K/8 x 3 dependent adds that bias reaches ~1e-5 relative at K = 512 (measured).
// The accumulator is therefore flushed every kChunkKB K blocks (K = 128): the MMA warp starts a
// fresh accumulator stage per chunk, and the epilogue warps add each finished chunk into fp32
// running sums in REGISTERS with round-to-nearest adds while the next chunk's MMAs run.
constexpr int kChunkKB = 4;

template <int BN, int MODE>
__global__ void __launch_bounds__(kTcThreads, 1) gemm_tc_kernel(const TcParams p) {
  constexpr int NSTAGE = (BN <= 128) ? 3 : 2;
  constexpr uint32_t kBTileBytes = BN * 32 * 4 * 2;
  constexpr uint32_t kStageBytes = kATileBytes + kBTileBytes;
  constexpr uint32_t kAccCols = 256;                       // column offset of the second accumulator stage
  // instruction descriptor: D = f32, A = B = tf32, both K-major, N = BN, M = 128
  constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) |
                              (static_cast<uint32_t>(TC_BM >> 4) << 24);
  extern __shared__ __align__(1024) uint8_t tc_smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(tc_smem + NSTAGE * kStageBytes);
  uint64_t* full = bars;                     // [NSTAGE]
  uint64_t* empty = bars + NSTAGE;           // [NSTAGE]
  uint64_t* tmem_full = bars + 2 * NSTAGE;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NSTAGE; ++s) mbar_init(&full[s], 1), mbar_init(&empty[s], 1);
    for (int s = 0; s < 2; ++s) mbar_init(&tmem_full[s], 1), mbar_init(&tmem_empty[s], kEpiWarps * 32);
    mbar_fence_init();
  }
  if (warp == kProducerWarp) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // barriers and TMEM are set up: wait for the kernels ahead in the stream before touching global memory, and let the
  // next kernel start its own set-up as soon as this grid's CTAs leave room
  pdl_wait();
  pdl_trigger();
  const uint32_t tmem_base = *tmem_slot;
  const int mn_tiles = p.m_tiles * p.n_tiles;
  const bool split = MODE == kTcPlain && p.ksplit > 1;
  const int num_tiles = split ? mn_tiles * p.ksplit : mn_tiles;

  if (warp == kProducerWarp) {
    // ===== producer =====
    if (lane == 0) {
      int stage = 0, issued = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mn = tile % mn_tiles, z = tile / mn_tiles;
        const int m_blk = mn % p.m_tiles, n_blk = mn / p.m_tiles;
        const int kb_lo = split ? z * p.kb_per : 0, kb_hi = split ? min(p.KB, kb_lo + p.kb_per) : p.KB;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (p.debug_noload && issued++ >= NSTAGE) {
            mbar_arrive(&full[stage]);
            if (++stage == NSTAGE) stage = 0, phase ^= 1;
            continue;
          }
          uint8_t* sa = tc_smem + stage * kStageBytes;
          mbar_arrive_expect_tx(&full[stage], kStageBytes);
          bulk_g2s(sa, p.A + (static_cast<size_t>(kb) * p.a_rg + static_cast<size_t>(m_blk) * (TC_BM / 8)) * 512,
                   kATileBytes, &full[stage]);
          bulk_g2s(sa + kATileBytes,
                   p.B + (static_cast<size_t>(kb) * p.b_rg + static_cast<size_t>(n_blk) * (BN / 8)) * 512, kBTileBytes,
                   &full[stage]);
          if (++stage == NSTAGE) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0, acc = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int kb_lo = split ? (tile / mn_tiles) * p.kb_per : 0;
        const int kb_hi = split ? min(p.KB, kb_lo + p.kb_per) : p.KB;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          const int in_chunk = (kb - kb_lo) % kChunkKB;
          if (in_chunk == 0) {                   // fresh accumulator stage for this K chunk
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            tc_fence_after();
          }
          const uint32_t d_tmem = tmem_base + acc * kAccCols;
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(tc_smem + stage * kStageBytes);
          const uint32_t sb = sa + kATileBytes;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const uint64_t a_hi = make_desc(sa + ks * 256), a_lo = make_desc(sa + 1024 + ks * 256);
            const uint64_t b_hi = make_desc(sb + ks * 256), b_lo = make_desc(sb + 1024 + ks * 256);
            umma_tf32(d_tmem, a_hi, b_lo, kIdesc, (in_chunk | ks) != 0);
            umma_tf32(d_tmem, a_lo, b_hi, kIdesc, 1);
            umma_tf32(d_tmem, a_hi, b_hi, kIdesc, 1);
          }
          umma_commit(&empty[stage]);            // smem slot free once these MMAs have read it
          if (++stage == NSTAGE) stage = 0, phase ^= 1;
          if (in_chunk == kChunkKB - 1 || kb == kb_hi - 1) {
            umma_commit(&tmem_full[acc]);        // chunk complete
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
          }
        }
      }
    }
  } else {
    // ===== epilogue (8 warps: TMEM lane quadrant = warp % 4, column half = warp / 4) =====
    constexpr bool kWide = MODE == kTcLayer || MODE == kTcPlain;   // thread owns half the tile's columns of its row
    constexpr int NG = kWide ? BN / 64 : 3;                  // 32-column groups owned by this thread
    const int quad = warp & 3, half = warp >> 2;
    uint32_t acc = 0, acc_phase = 0;
    const int row_in_tile = quad * 32 + lane;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int mn = tile % mn_tiles, z = tile / mn_tiles;
      const int m_blk = mn % p.m_tiles, n_blk = mn / p.m_tiles;
      const int kb_lo = split ? z * p.kb_per : 0, kb_hi = split ? min(p.KB, kb_lo + p.kb_per) : p.KB;
      const int num_chunks = (kb_hi - kb_lo + kChunkKB - 1) / kChunkKB;
      const int r = m_blk * TC_BM + row_in_tile;
      float sum[NG][32];
#pragma unroll
      for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int j = 0; j < 32; ++j) sum[g][j] = 0.f;
      // plain epilogue: the ReLU gate of this thread's row and columns as one bit per column, fetched NOW -- while the
      // MMAs of the tile run -- instead of after the accumulator arrives
      uint32_t gate_bits[NG];
#pragma unroll
      for (int g = 0; g < NG; ++g) gate_bits[g] = 0xffffffffu;
      if (MODE == kTcPlain && p.gate != nullptr && !split && r < p.rows) {
#pragma unroll
        for (int g = 0; g < NG; ++g) {
          const int n0 = n_blk * BN + half * (BN / 2) + g * 32;
          const float* gt = p.gate + static_cast<long long>(r) * p.ldg + n0;
          uint32_t bits = 0;
          if (n0 + 31 < p.cols && ((reinterpret_cast<uintptr_t>(gt) & 15u) == 0)) {
#pragma unroll
            for (int j4 = 0; j4 < 32; j4 += 4) {
              const float4 t = *reinterpret_cast<const float4*>(gt + j4);
              bits |= (t.x > 0.f ? 1u : 0u) << j4;
              bits |= (t.y > 0.f ? 1u : 0u) << (j4 + 1);
              bits |= (t.z > 0.f ? 1u : 0u) << (j4 + 2);
              bits |= (t.w > 0.f ? 1u : 0u) << (j4 + 3);
            }
          } else {
            for (int j = 0; j < 32; ++j)
              if (n0 + j < p.cols && gt[j] > 0.f) bits |= 1u << j;
          }
          gate_bits[g] = bits;
        }
      }
      for (int chunk = 0; chunk < num_chunks; ++chunk) {
        mbar_wait(&tmem_full[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * kAccCols;
#pragma unroll
        for (int g = 0; g < NG; ++g) {
          const int col = kWide ? half * (BN / 2) + g * 32 : g * 64 + half * 32;
          float v[32];
          tmem_ld32(taddr + col, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) sum[g][j] = __fadd_rn(sum[g][j], v[j]);
        }
        tc_fence_before();
        mbar_arrive(&tmem_empty[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      const bool alt = p.row_set != nullptr && r < p.rows && p.row_set[r] != 0;
      if (MODE == kTcPlain) {
        if (r < p.rows) {
          // split-K: raw sums of this K range; gate / scale / accumulate happen in the reduce stage
          const float* cs = split ? nullptr : alt ? p.cscale_alt : p.cscale;
          const int cinv = alt ? p.cinv_alt : p.cinv;
          const int accumulate = split ? 0 : p.accumulate;
          float* crow = split ? p.partial + (static_cast<size_t>(z) * p.rows + r) * p.cols
                              : p.C + static_cast<long long>(r) * p.ldc;
#pragma unroll
          for (int g = 0; g < NG; ++g) {
            const int n0 = n_blk * BN + half * (BN / 2) + g * 32;
            float* dst = crow + n0;
#pragma unroll
            for (int j4 = 0; j4 < 32; j4 += 4) {
              float o[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int n = n0 + j4 + j;
                float val = sum[g][j4 + j];
                if (n < p.cols) {
                  if (!((gate_bits[g] >> (j4 + j)) & 1u)) val = 0.f;
                  if (cs != nullptr) val = __fmul_rn(val, cinv ? __fsub_rn(1.f, cs[n]) : cs[n]);
                  if (accumulate) val = __fadd_rn(dst[j4 + j], val);
                }
                o[j] = val;
              }
              if (n0 + j4 + 3 < p.cols && ((reinterpret_cast<uintptr_t>(dst + j4) & 15u) == 0)) {
                *reinterpret_cast<float4*>(dst + j4) = make_float4(o[0], o[1], o[2], o[3]);
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (n0 + j4 + j < p.cols) dst[j4 + j] = o[j];
              }
              if (p.out_ts != nullptr && !split) {
                // the same values as the A operand (TS32) of the product that consumes C next: K block n0 / 32 of its
                // operand, columns past `cols` are this product's zero padding (o = 0 there)
                const int kb = n0 >> 5;
                if (kb < p.out_kb) {
                  float h[4], l[4];
#pragma unroll
                  for (int j = 0; j < 4; ++j) h[j] = tf32_hi(o[j]), l[j] = tf32_lo(o[j], h[j]);
                  float* ts = p.out_ts + ((static_cast<size_t>(kb) * p.out_rg + (r >> 3)) * 2) * 256 + (r & 7) * 4 + (j4 >> 2) * 32;
                  *reinterpret_cast<float4*>(ts) = make_float4(h[0], h[1], h[2], h[3]);
                  *reinterpret_cast<float4*>(ts + 256) = make_float4(l[0], l[1], l[2], l[3]);
                }
              }
            }
          }
        }
      } else if (MODE == kTcLayer) {
        const int rg = r >> 3, r8 = r & 7;
        const float* bias = alt ? p.bias_alt : p.bias;
#pragma unroll
        for (int g = 0; g < NG; ++g) {
          const int kb = n_blk * (BN / 32) + half * (BN / 64) + g;     // K block of the NEXT layer
          if (kb < p.out_kb) {
            const int n0 = kb * 32;
            float* dst = p.out_ts + ((static_cast<size_t>(kb) * p.out_rg + rg) * 2) * 256 + r8 * 4;
#pragma unroll
            for (int kc = 0; kc < 8; ++kc) {
              float h[4], l[4], o[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int n = n0 + kc * 4 + j;
                const float b = n < p.cols ? __ldg(bias + n) : 0.f;
                o[j] = fmaxf(__fadd_rn(sum[g][kc * 4 + j], b), 0.f);
                h[j] = tf32_hi(o[j]);
                l[j] = tf32_lo(o[j], h[j]);
              }
              *reinterpret_cast<float4*>(dst + kc * 32) = make_float4(h[0], h[1], h[2], h[3]);
              *reinterpret_cast<float4*>(dst + 256 + kc * 32) = make_float4(l[0], l[1], l[2], l[3]);
              if (p.out_rm != nullptr && r < p.rows) {
                float* rm = p.out_rm + static_cast<size_t>(r) * p.cols + n0 + kc * 4;
                if (n0 + kc * 4 + 3 < p.cols && (reinterpret_cast<uintptr_t>(rm) & 15u) == 0) {
                  *reinterpret_cast<float4*>(rm) = make_float4(o[0], o[1], o[2], o[3]);
                } else {
#pragma unroll
                  for (int j = 0; j < 4; ++j)
                    if (n0 + kc * 4 + j < p.cols) rm[j] = o[j];
                }
              }
            }
          }
        }
      } else {
        // heads: sum[0] = S, sum[1] = T, sum[2] = Q pre-activations of d = 64*n_blk + 32*half + j
        const bool rok = r < p.rows;
        const float heps = __fmul_rn(0.5f, p.eps);
        const int dir = alt ? p.dir_alt : p.dir;
        float ld = 0.f;
        bool bad = false;
        // per-output constants of this thread's 32 outputs: lane l fetches those of output d = 64 n_blk + 32 half + l
        // once (two expf per thread instead of two per element); the loop below reads them by shuffle
        const int dl = n_blk * 64 + half * 32 + lane;
        const bool dl_ok = dl < p.cols;
        const float bs_l = dl_ok ? __ldg(p.b_s + dl) : 0.f, bt_l = dl_ok ? __ldg(p.b_tr + dl) : 0.f;
        const float bq_l = dl_ok ? __ldg(p.b_q + dl) : 0.f;
        const float es_l = dl_ok ? expf(__ldg(p.coeff_s + dl)) : 1.f, eq_l = dl_ok ? expf(__ldg(p.coeff_q + dl)) : 1.f;
        float m0_l = 0.f, m1_l = 0.f;
        if (MODE == kTcHeadsPosition && dl_ok) {
          m0_l = __ldg(p.umask + dl);
          if (p.uinvert) m0_l = __fsub_rn(1.f, m0_l);
          if (p.row_set != nullptr) {
            m1_l = __ldg(p.umask_alt + dl);
            if (p.uinvert_alt) m1_l = __fsub_rn(1.f, m1_l);
          }
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int d0 = n_blk * 64 + half * 32 + c * 16;
          if (d0 < p.cols) {              // warp-uniform; cols is a multiple of 16 on this path
            const size_t at = static_cast<size_t>(r) * p.cols + d0;
            float a0[16], a1[16], o0[16], o1[16], o2[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) a0[j] = a1[j] = 0.f;
            if (!rok) {
            } else if (MODE == kTcHeadsMomentum) {
#pragma unroll
              for (int j = 0; j < 16; j += 4) {
                *reinterpret_cast<float4*>(a0 + j) = *reinterpret_cast<const float4*>(p.uv + at + j);
                *reinterpret_cast<float4*>(a1 + j) = *reinterpret_cast<const float4*>(p.ug + at + j);
              }
            } else if (MODE == kTcHeadsPosition) {
#pragma unroll
              for (int j = 0; j < 16; j += 4) {
                *reinterpret_cast<float4*>(a0 + j) = *reinterpret_cast<const float4*>(p.ux + at + j);
                *reinterpret_cast<float4*>(a1 + j) = *reinterpret_cast<const float4*>(p.uv + at + j);
              }
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int src = c * 16 + j;
              const float bs = __shfl_sync(0xffffffffu, bs_l, src), es = __shfl_sync(0xffffffffu, es_l, src);
              const float bt = __shfl_sync(0xffffffffu, bt_l, src), bq = __shfl_sync(0xffffffffu, bq_l, src);
              const float eq = __shfl_sync(0xffffffffu, eq_l, src);
              const float S = __fmul_rn(tanhf(__fadd_rn(sum[0][c * 16 + j], bs)), es);
              const float Tr = __fadd_rn(sum[1][c * 16 + j], bt);
              float Q = __fadd_rn(sum[2][c * 16 + j], bq);
              if (p.tanh_on_q) Q = tanhf(Q);
              Q = __fmul_rn(Q, eq);
              if (MODE == kTcHeadsStq) {
                o0[j] = S, o1[j] = Tr, o2[j] = Q;
              } else if (MODE == kTcHeadsMomentum) {
                o0[j] = dir > 0 ? momentum_elem<1>(a0[j], a1[j], S, Tr, Q, p.eps, heps, ld, bad)
                                : momentum_elem<-1>(a0[j], a1[j], S, Tr, Q, p.eps, heps, ld, bad);
              } else {
                const float mf = __shfl_sync(0xffffffffu, m0_l, src), ma = __shfl_sync(0xffffffffu, m1_l, src);
                const float m = alt ? ma : mf;
                o0[j] = dir > 0 ? position_elem<1>(a0[j], a1[j], S, Tr, Q, m, p.eps, ld, bad)
                                : position_elem<-1>(a0[j], a1[j], S, Tr, Q, m, p.eps, ld, bad);
              }
            }
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              if (!rok) {
              } else if (MODE == kTcHeadsStq) {
                *reinterpret_cast<float4*>(p.S + at + j) = *reinterpret_cast<float4*>(o0 + j);
                *reinterpret_cast<float4*>(p.T + at + j) = *reinterpret_cast<float4*>(o1 + j);
                *reinterpret_cast<float4*>(p.Q + at + j) = *reinterpret_cast<float4*>(o2 + j);
              } else {
                *reinterpret_cast<float4*>(p.uout + at + j) = *reinterpret_cast<float4*>(o0 + j);
              }
            }
          }
        }
        if (MODE != kTcHeadsStq) {
          // one partial per (64-wide head tile, column half): [2*n_blk + half][rows]
          if (rok && p.ld_part) p.ld_part[static_cast<size_t>(2 * n_blk + half) * p.rows + r] = ld;
          flag_nonfinite(p.flag, bad);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kProducerWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------ format kernels
// row-major fp32 [rows, K] (leading dimension ld, optional 0/1 mask row) -> TS32 K blocks
// [kb0, kb0 + ceil(K/32)) of a buffer with `rg_total` row groups; rows >= `rows` and k >= K are zero.
__device__ __forceinline__ void ts32_from_rowmajor(const float* __restrict__ src, int ld, int rows, int K,
                                                   const float* __restrict__ mask, int invert,
                                                   const float* __restrict__ mask_alt, int invert_alt,
                                                   const uint8_t* __restrict__ row_set, float* dst, int rg_total, int kb0,
                                                   int nkb) {
  // one thread per (kb, rg, kc, r8): consecutive threads write consecutive 16 B
  const size_t total = static_cast<size_t>(nkb) * rg_total * 64;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int r8 = static_cast<int>(i & 7), kc = static_cast<int>((i >> 3) & 7);
    const size_t blk = i >> 6;
    const int rg = static_cast<int>(blk % rg_total), kb = static_cast<int>(blk / rg_total);
    const int r = rg * 8 + r8, k = kb * 32 + kc * 4;
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    if (r < rows) {
      const float* s = src + static_cast<size_t>(r) * ld + k;
      if (k + 3 < K && ((reinterpret_cast<uintptr_t>(s) & 15u) == 0)) {
        const float4 t = *reinterpret_cast<const float4*>(s);
        a[0] = t.x, a[1] = t.y, a[2] = t.z, a[3] = t.w;
      } else {
        for (int j = 0; j < 4; ++j)
          if (k + j < K) a[j] = s[j];
      }
      const bool alt = row_set != nullptr && row_set[r] != 0;
      const float* mrow = alt ? mask_alt : mask;
      if (mrow) {
        const int inv = alt ? invert_alt : invert;
        for (int j = 0; j < 4; ++j)
          if (k + j < K) {
            const float m = mrow[k + j];
            a[j] = __fmul_rn(inv ? __fsub_rn(1.f, m) : m, a[j]);
          }
      }
    }
    float h[4], l[4];
    for (int j = 0; j < 4; ++j) h[j] = tf32_hi(a[j]), l[j] = tf32_lo(a[j], h[j]);
    float* d = dst + ((static_cast<size_t>(kb0 + kb) * rg_total + rg) * 2) * 256 + kc * 32 + r8 * 4;
    *reinterpret_cast<float4*>(d) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(d + 256) = make_float4(l[0], l[1], l[2], l[3]);
  }
}
__global__ void __launch_bounds__(256) ts32_from_rowmajor_kernel(const float* __restrict__ src, int ld, int rows, int K,
                                                                 const float* __restrict__ mask, int invert,
                                                                 const float* __restrict__ mask_alt, int invert_alt,
                                                                 const uint8_t* __restrict__ row_set,
                                                                 float* dst, int rg_total, int kb0, int nkb) {
  ts32_from_rowmajor(src, ld, rows, K, mask, invert, mask_alt, invert_alt, row_set, dst, rg_total, kb0, nkb);
}

// Everything a network call needs before its first GEMM, in ONE launch (it was four: at config 3 each was ~4 us of
// launch latency in a 136 us call): blockIdx.y = 0 / 1: the two inputs -> TS32 (mask folded into the second),
// blockIdx.y = 2: the effective first-layer bias of both time-embedding sets.
struct NetPrologueParams {
  const float *in0, *in1;
  int K, N;
  const float* mask;
  int invert;
  const float* mask_alt;
  int invert_alt;
  const uint8_t* row_set;
  float* acat;
  int rg, kbK;
  const float *b0, *b1, *b2, *wt;
  float tau0, tau1, tau0_alt, tau1_alt;
  int H, h_rows;
  float* bias_eff;
};
__global__ void __launch_bounds__(256) net_prologue_kernel(const NetPrologueParams q) {
  pdl_wait();
  pdl_trigger();
  if (blockIdx.y == 0) {
    ts32_from_rowmajor(q.in0, q.K, q.N, q.K, nullptr, 0, nullptr, 0, nullptr, q.acat, q.rg, 0, q.kbK);
  } else if (blockIdx.y == 1) {
    ts32_from_rowmajor(q.in1, q.K, q.N, q.K, q.mask, q.invert, q.mask_alt, q.invert_alt, q.row_set, q.acat, q.rg, q.kbK,
                       q.kbK);
  } else {
    // bias_eff[n] = b0[n] + b1[n] + (tau0*wt[0][n] + tau1*wt[1][n] + b2[n]), the op order of bias_eff_kernel
    const int sets = q.row_set != nullptr ? 2 : 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < sets * q.H; i += gridDim.x * blockDim.x) {
      const int set = i / q.H, n = i - set * q.H;
      const float t0 = set ? q.tau0_alt : q.tau0, t1 = set ? q.tau1_alt : q.tau1;
      const float tt = __fadd_rn(__fadd_rn(__fmul_rn(t0, q.wt[n]), __fmul_rn(t1, q.wt[q.H + n])), q.b2[n]);
      q.bias_eff[set * q.h_rows + n] = __fadd_rn(__fadd_rn(q.b0[n], q.b1[n]), tt);
    }
  }
}

// Keras kernel W [K, cols] (row-major) -> TS32 of W^T: packed row `prow(n)` = output feature n, K blocks
// [kb0, kb0 + ceil(K/32)).  heads != 0: n = output d of head `head`, packed row = (d/64)*192 + head*64 + d%64.
__global__ void __launch_bounds__(256) ts32_pack_weight_kernel(const float* __restrict__ W, int K, int cols, float* dst,
                                                               int rg_total, int kb0, int heads, int head) {
  const int nkb = (K + 31) / 32;
  const size_t total = static_cast<size_t>(nkb) * 32 * cols;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int n = static_cast<int>(i % cols);
    const int k = static_cast<int>(i / cols);
    const float a = k < K ? W[static_cast<size_t>(k) * cols + n] : 0.f;
    const int pr = heads ? (n / 64) * 192 + head * 64 + (n % 64) : n;
    const int kb = k / 32, kc = (k % 32) / 4, k4 = k % 4;
    float* d = dst + ((static_cast<size_t>(kb0 + kb) * rg_total + pr / 8) * 2) * 256 + kc * 32 + (pr % 8) * 4 + k4;
    const float h = tf32_hi(a);
    d[0] = h;
    d[256] = tf32_lo(a, h);
  }
}

// TS32 (hi + lo) -> row-major fp32 (diagnostics)
__global__ void __launch_bounds__(256) ts32_to_rowmajor_kernel(const float* __restrict__ src, int rg_total, int rows,
                                                               int K, float* dst) {
  const size_t total = static_cast<size_t>(rows) * K;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % K), r = static_cast<int>(i / K);
    const float* s = src + ((static_cast<size_t>(k / 32) * rg_total + r / 8) * 2) * 256 + ((k % 32) / 4) * 32 +
                     (r % 8) * 4 + k % 4;
    dst[i] = __fadd_rn(s[0], s[256]);
  }
}

__global__ void logdet_reduce_tc_kernel(const float* __restrict__ part, int tiles, int N, float* logdet) {
  pdl_wait();
  pdl_trigger();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int j = 0; j < tiles; ++j) s = __fadd_rn(s, part[static_cast<size_t>(j) * N + n]);
  logdet[n] = __fadd_rn(logdet[n], s);
}

inline int ew_grid(size_t total) { return static_cast<int>(std::min<size_t>((total + 255) / 256, 148 * 32)); }

template <int BN, int MODE>
int launch_tc(const TcParams& p, cudaStream_t st) {
  constexpr int NSTAGE = (BN <= 128) ? 3 : 2;
  constexpr size_t smem = static_cast<size_t>(NSTAGE) * (kATileBytes + BN * 256) + (2 * NSTAGE + 4) * 8 + 16;
  auto kern = gemm_tc_kernel<BN, MODE>;
  // per device: the opt-in shared-memory size and the SM count (one process may drive several GPUs)
  constexpr int kMaxDevices = 64;
  static bool opted[kMaxDevices] = {};
  static int sm_count[kMaxDevices] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  const int slot = dev < kMaxDevices ? dev : 0;
  if (dev >= kMaxDevices || !opted[slot]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(gemm_tc_kernel, %zu B): %s", smem, cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&sm_count[slot], cudaDevAttrMultiProcessorCount, dev);
    opted[slot] = dev < kMaxDevices;
  }
  const int sms = sm_count[slot];
  const int tiles = p.m_tiles * p.n_tiles * (MODE == kTcPlain && p.ksplit > 1 ? p.ksplit : 1);
  if (tiles == 0) return L2HMC_OK;
  launch_pdl(kern, dim3(std::min(tiles, sms)), dim3(kTcThreads), smem, st, p);
  return check_launch("gemm_tc_kernel");
}

}  // namespace

// ------------------------------------------------------------------ sizes shared with the pack / workspace API
TcDims tc_dims(int K, int H, int D, int N) {
  TcDims d;
  d.kbK = (K + 31) / 32;
  d.kbH = (H + 31) / 32;
  d.h_rows = (H + 255) / 256 * 256;            // packed rows of W1^T / W2^T (covers BN = 128 and 256)
  d.d_tiles = (D + 63) / 64;
  d.rows_p = (N + 127) / 128 * 128;
  d.w1_floats = static_cast<size_t>(2 * d.kbK) * d.h_rows * 64;     // rows * kb * 32 k * 2 parts
  d.w2_floats = static_cast<size_t>(d.kbH) * d.h_rows * 64;
  d.w3_floats = static_cast<size_t>(d.kbH) * (d.d_tiles * 192) * 64;
  d.acat_floats = static_cast<size_t>(2 * d.kbK) * d.rows_p * 64;
  d.h_floats = static_cast<size_t>(d.kbH) * d.rows_p * 64;
  return d;
}

// structural requirements of the tcgen05 path: packed weights, D a multiple of the head epilogue's
// 16-column chunks (16-byte aligned rows for the update operands); rows / K / H are padded
bool tc_supported(const l2hmc_net_weights* w, const l2hmc_net_call* c) {
  (void)c;
  return w->packed != nullptr && (w->D % 16) == 0;
}

// AUTO: dense-contraction shapes only (north star) -- enough chains to fill 128-row MMA tiles and to beat the chunked
// FFMA layers (measured, profiles/r2_small_ops_sweep.log: a tcgen05 net call costs ~50 us whatever the batch up to
// ~2048 rows; the FFMA call costs 28 us at 256 x 512 hidden, 75 us at 1024 x 512): from N x H >= 2^19, at least 256.
int tc_min_rows(int K, int H, int D) {
  if (!(K >= 64 && H >= 128 && D >= 64 && D % 16 == 0)) return INT_MAX;
  return std::max(256, (524288 + H - 1) / H);
}
bool tc_eligible(const l2hmc_net_weights* w, const l2hmc_net_call* c) {
  return tc_supported(w, c) && c->N >= tc_min_rows(w->K, w->H, w->D);
}

int net_run_tc(const l2hmc_net_weights* w, const l2hmc_net_call* c, cudaStream_t st) {
  const int N = c->N, K = w->K, H = w->H, D = w->D;
  const TcDims d = tc_dims(K, H, D, N);
  const float* packed = static_cast<const float*>(w->packed);
  const float* W1 = packed;
  const float* W2 = W1 + d.w1_floats;
  const float* W3 = W2 + d.w2_floats;
  float* ws = static_cast<float*>(c->workspace);
  float* acat = ws;
  float* h1 = acat + d.acat_floats;
  float* h2 = h1 + d.h_floats;
  float* bias_eff = h2 + d.h_floats;
  float* ld_part = bias_eff + 2 * d.h_rows;       // [2][h_rows]: second row for the *_alt time embedding
  const int rg = d.rows_p / 8;

  // network inputs -> TS32 (mask folded into the second input)
  NetPrologueParams q{c->in0, c->in1, K, N, c->in1_mask, c->in1_mask_invert, c->in1_mask_alt, c->in1_mask_invert_alt,
                      c->row_set, acat, rg, d.kbK, w->b_in0, w->b_in1, w->b_t, w->w_t, c->tau_cos, c->tau_sin,
                      c->tau_cos_alt, c->tau_sin_alt, H, d.h_rows, bias_eff};
  launch_pdl(net_prologue_kernel, dim3(ew_grid(static_cast<size_t>(d.kbK) * rg * 64), 3), dim3(256), 0, st, q);
  int rc = check_launch("net_prologue_kernel");
  if (rc) return rc;

  static const int debug_noload = getenv("L2HMC_TC_DEBUG_NOLOAD") ? atoi(getenv("L2HMC_TC_DEBUG_NOLOAD")) : 0;
  const int m_tiles = d.rows_p / 128;
  // few tiles: halve the N tile so more SMs take part
  const bool narrow = m_tiles * ((H + 255) / 256) < 120;
  TcParams p{};
  p.A = acat, p.a_rg = rg, p.B = W1, p.b_rg = d.h_rows / 8, p.KB = 2 * d.kbK;
  p.m_tiles = m_tiles, p.n_tiles = narrow ? (H + 127) / 128 : (H + 255) / 256;
  p.debug_noload = debug_noload;
  p.rows = N, p.cols = H, p.bias = bias_eff, p.bias_alt = bias_eff + d.h_rows, p.row_set = c->row_set;
  p.out_ts = h1, p.out_rg = rg, p.out_kb = d.kbH, p.out_rm = c->h1_out;
  rc = narrow ? launch_tc<128, kTcLayer>(p, st) : launch_tc<256, kTcLayer>(p, st);
  if (rc) return rc;

  p.A = h1, p.B = W2, p.KB = d.kbH, p.bias = w->b_h, p.bias_alt = nullptr, p.row_set = nullptr, p.out_ts = h2;
  p.out_rm = c->h2_out;
  rc = narrow ? launch_tc<128, kTcLayer>(p, st) : launch_tc<256, kTcLayer>(p, st);
  if (rc) return rc;

  p = TcParams{};
  p.debug_noload = debug_noload;
  p.A = h2, p.a_rg = rg, p.B = W3, p.b_rg = d.d_tiles * 192 / 8, p.KB = d.kbH;
  p.m_tiles = m_tiles, p.n_tiles = d.d_tiles;
  p.rows = N, p.cols = D;
  p.b_s = w->b_s, p.b_tr = w->b_tr, p.b_q = w->b_q, p.coeff_s = w->coeff_s, p.coeff_q = w->coeff_q;
  p.tanh_on_q = w->tanh_on_q;
  p.S = c->S, p.T = c->T, p.Q = c->Q;
  p.ux = c->x, p.uv = c->v, p.ug = c->g, p.umask = c->mask, p.uinvert = c->mask_invert;
  p.eps = c->eps, p.dir = c->dir, p.uout = c->out, p.flag = c->nonfinite_flag;
  p.row_set = c->row_set, p.umask_alt = c->mask_alt, p.uinvert_alt = c->mask_invert_alt, p.dir_alt = c->dir_alt;
  p.ld_part = c->logdet ? ld_part : nullptr;
  if (c->mode == L2HMC_NET_STQ) rc = launch_tc<192, kTcHeadsStq>(p, st);
  else if (c->mode == L2HMC_NET_MOMENTUM) rc = launch_tc<192, kTcHeadsMomentum>(p, st);
  else rc = launch_tc<192, kTcHeadsPosition>(p, st);
  if (rc) return rc;
  if (c->mode != L2HMC_NET_STQ && c->logdet) {
    launch_pdl(logdet_reduce_tc_kernel, dim3((N + 255) / 256), dim3(256), 0, st, ld_part, 2 * d.d_tiles, N, c->logdet);
    rc = check_launch("logdet_reduce_tc_kernel");
  }
  // training tape (h1_out / h2_out): the layer epilogues above also wrote the hidden activations row-major fp32
  return rc;
}

size_t tc_workspace_floats(int K, int H, int D, int N) {
  const TcDims d = tc_dims(K, H, D, N);
  return d.acat_floats + 2 * d.h_floats + 2 * d.h_rows + 2 * static_cast<size_t>(d.d_tiles) * d.rows_p;
}

}  // namespace l2hmc

using namespace l2hmc;

extern "C" size_t l2hmc_net_pack_bytes(int K, int H, int D) {
  if (K <= 0 || H <= 0 || D <= 0) return 0;
  const TcDims d = tc_dims(K, H, D, 128);
  return sizeof(float) * (d.w1_floats + d.w2_floats + d.w3_floats);
}

extern "C" int l2hmc_net_pack(const l2hmc_net_weights* w, void* packed, size_t packed_bytes, void* stream) {
  L2HMC_REQUIRE(w && packed, "l2hmc_net_pack: null argument");
  L2HMC_REQUIRE(w->K > 0 && w->H > 0 && w->D > 0, "l2hmc_net_pack: bad sizes");
  L2HMC_REQUIRE(w->w_in0 && w->w_in1 && w->w_h && w->w_s && w->w_tr && w->w_q, "l2hmc_net_pack: null weight pointer");
  L2HMC_REQUIRE(packed_bytes >= l2hmc_net_pack_bytes(w->K, w->H, w->D), "l2hmc_net_pack: buffer too small");
  L2HMC_REQUIRE(aligned16(packed), "l2hmc_net_pack: buffer must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const TcDims d = tc_dims(w->K, w->H, w->D, 128);
  float* W1 = static_cast<float*>(packed);
  float* W2 = W1 + d.w1_floats;
  float* W3 = W2 + d.w2_floats;
  cudaError_t e = cudaMemsetAsync(packed, 0, l2hmc_net_pack_bytes(w->K, w->H, w->D), st);   // zero padding rows / k
  if (e != cudaSuccess) {
    set_error("l2hmc_net_pack: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return L2HMC_ERR_CUDA;
  }
  const int K = w->K, H = w->H, D = w->D;
  const size_t t1 = static_cast<size_t>(d.kbK) * 32 * H, t2 = static_cast<size_t>(d.kbH) * 32 * H;
  const size_t t3 = static_cast<size_t>(d.kbH) * 32 * D;
  ts32_pack_weight_kernel<<<ew_grid(t1), 256, 0, st>>>(w->w_in0, K, H, W1, d.h_rows / 8, 0, 0, 0);
  ts32_pack_weight_kernel<<<ew_grid(t1), 256, 0, st>>>(w->w_in1, K, H, W1, d.h_rows / 8, d.kbK, 0, 0);
  ts32_pack_weight_kernel<<<ew_grid(t2), 256, 0, st>>>(w->w_h, H, H, W2, d.h_rows / 8, 0, 0, 0);
  ts32_pack_weight_kernel<<<ew_grid(t3), 256, 0, st>>>(w->w_s, H, D, W3, d.d_tiles * 192 / 8, 0, 1, 0);
  ts32_pack_weight_kernel<<<ew_grid(t3), 256, 0, st>>>(w->w_tr, H, D, W3, d.d_tiles * 192 / 8, 0, 1, 1);
  ts32_pack_weight_kernel<<<ew_grid(t3), 256, 0, st>>>(w->w_q, H, D, W3, d.d_tiles * 192 / 8, 0, 1, 2);
  return check_launch("ts32_pack_weight_kernel", 6);
}

// ------------------------------------------------------------------ general tcgen05 GEMM (training backward)
namespace l2hmc {
namespace {

// strided fp32 operand -> TS32: element (r, k) = src[r * rs + k * cs]; rows >= `rows` and k >= K are zero; K blocks
// [kb0, kb0 + nkb) of a buffer with rg_total row groups.  Serves A (rows = output rows) and B (rows = output columns)
// in NN / NT / TN forms without a transposed copy.
// One launch converts every operand segment of a product (blockIdx.y = job): up to 3 segments x (A, B).
struct Ts32Job {
  const float* src;
  long long rs, cs;
  int rows, K;
  float* dst;
  int rg_total, kb0, nkb;
};
struct Ts32Jobs {
  Ts32Job j[6];
};
__global__ void __launch_bounds__(256) ts32_from_strided_kernel(const Ts32Jobs jobs) {
  pdl_wait();
  pdl_trigger();
  const Ts32Job& jb = jobs.j[blockIdx.y];
  const float* __restrict__ src = jb.src;
  const long long rs = jb.rs, cs = jb.cs;
  const int rows = jb.rows, K = jb.K, rg_total = jb.rg_total, kb0 = jb.kb0, nkb = jb.nkb;
  float* dst = jb.dst;
  const size_t total = static_cast<size_t>(nkb) * rg_total * 64;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int r8 = static_cast<int>(i & 7), kc = static_cast<int>((i >> 3) & 7);
    const size_t blk = i >> 6;
    const int rg = static_cast<int>(blk % rg_total), kb = static_cast<int>(blk / rg_total);
    const int r = rg * 8 + r8, k = kb * 32 + kc * 4;
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    if (r < rows) {
      const float* s = src + static_cast<long long>(r) * rs + static_cast<long long>(k) * cs;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (k + j < K) a[j] = s[j * cs];
    }
    float h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = tf32_hi(a[j]), l[j] = tf32_lo(a[j], h[j]);
    float* d = dst + ((static_cast<size_t>(kb0 + kb) * rg_total + rg) * 2) * 256 + kc * 32 + r8 * 4;
    *reinterpret_cast<float4*>(d) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(d + 256) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

struct TcGemmDims {
  int rows_p, n_p, kb;
  size_t a_floats, b_floats;
  bool narrow;          // 128-wide N tiles (few tiles: more SMs take part)
  int n_tiles;
  int ksplit, kb_per;   // split-K: ksplit K ranges of kb_per K blocks each (the last one shorter), 1 = off
  size_t partial_floats;
};
constexpr int kTcSplitTargetCtas = 148;   // B200 SM count; a fixed constant so that results do not depend on the device
constexpr int kTcSplitMinKB = 32;         // at least K = 1024 per range
TcGemmDims tc_gemm_dims(const l2hmc_gemm_args* g) {
  TcGemmDims d{};
  d.rows_p = (g->M + 127) / 128 * 128;
  d.n_p = (g->N + 255) / 256 * 256;
  for (int s = 0; s < g->nseg; ++s) d.kb += (g->seg[s].K + 31) / 32;
  d.a_floats = static_cast<size_t>(d.kb) * d.rows_p * 64;
  d.b_floats = static_cast<size_t>(d.kb) * d.n_p * 64;
  const int m_tiles = d.rows_p / 128;
  d.narrow = m_tiles * ((g->N + 255) / 256) < 120;
  d.n_tiles = d.narrow ? (g->N + 127) / 128 : (g->N + 255) / 256;
  // long-K products with few output tiles (the weight gradients over the whole tape): split K over the idle SMs
  const int tiles = m_tiles * d.n_tiles;
  int ks = tiles > 0 ? std::min(kTcSplitTargetCtas / tiles, d.kb / kTcSplitMinKB) : 1;
  d.ksplit = 1, d.kb_per = d.kb;
  if (ks >= 2) {
    const int per = ((d.kb + ks - 1) / ks + kChunkKB - 1) / kChunkKB * kChunkKB;
    ks = (d.kb + per - 1) / per;          // no empty range
    if (ks >= 2) d.ksplit = ks, d.kb_per = per;
  }
  d.partial_floats = d.ksplit > 1 ? static_cast<size_t>(d.ksplit) * g->M * g->N : 0;
  return d;
}

}  // namespace
}  // namespace l2hmc

extern "C" int l2hmc_net_tc_min_rows(int K, int H, int D) { return l2hmc::tc_min_rows(K, H, D); }

extern "C" size_t l2hmc_tc_gemm_workspace_bytes(const l2hmc_gemm_args* g) {
  if (!g || g->nseg < 1 || g->nseg > 3 || g->M <= 0 || g->N <= 0) return 0;
  const TcGemmDims d = tc_gemm_dims(g);
  return sizeof(float) * (d.a_floats + d.b_floats + d.partial_floats) + 256;
}

extern "C" int l2hmc_tc_gemm(const l2hmc_gemm_args* g, void* workspace, size_t workspace_bytes, void* stream) {
  return l2hmc::tc_gemm_chain(g, nullptr, nullptr, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

// Floats of the TS32 form of an [M, K] A operand (what tc_gemm_chain's a_ts / out_ts point to).
size_t l2hmc::tc_gemm_ts_floats(int M, int K) {
  return static_cast<size_t>((K + 31) / 32) * ((M + 127) / 128 * 128) * 64;
}

// l2hmc_tc_gemm for a CHAIN of products (the backward pass of the dense trunk: d z2 -> d z1 -> input gradients):
// `out_ts` (nullable) also receives C in the tiled TF32 hi/lo form of an A operand [M, N], written by the epilogue;
// `a_ts` (nullable, one K segment only) is segment 0's A already in that form -- the product that made it saved this
// one an operand-conversion pass over A.  out_ts needs an unsplit product (every per-call product is).
int l2hmc::tc_gemm_chain(const l2hmc_gemm_args* g, const float* a_ts_in, float* out_ts, void* workspace,
                         size_t workspace_bytes, cudaStream_t st) {
  L2HMC_REQUIRE(g, "l2hmc_tc_gemm: null arguments");
  L2HMC_REQUIRE(g->M >= 0 && g->N >= 0 && g->nseg >= 1 && g->nseg <= 3, "l2hmc_tc_gemm: bad sizes M=%d N=%d nseg=%d", g->M,
                g->N, g->nseg);
  if (g->M == 0 || g->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(g->C, "l2hmc_tc_gemm: null C");
  for (int s = 0; s < g->nseg; ++s)
    L2HMC_REQUIRE(g->seg[s].A && g->seg[s].B && g->seg[s].K >= 0, "l2hmc_tc_gemm: bad segment %d", s);
  if (g->row_set) L2HMC_REQUIRE(!g->col_scale == !g->col_scale_alt, "l2hmc_tc_gemm: col_scale and col_scale_alt go together");
  L2HMC_REQUIRE(workspace && workspace_bytes >= l2hmc_tc_gemm_workspace_bytes(g) && aligned16(workspace),
                "l2hmc_tc_gemm: workspace too small or misaligned (%zu < %zu bytes)", workspace_bytes,
                l2hmc_tc_gemm_workspace_bytes(g));
  const TcGemmDims d = tc_gemm_dims(g);
  L2HMC_REQUIRE(a_ts_in == nullptr || g->nseg == 1, "tc_gemm_chain: a pre-tiled A operand needs a single K segment");
  L2HMC_REQUIRE(out_ts == nullptr || d.ksplit == 1, "tc_gemm_chain: the tiled output needs an unsplit product");
  float* a_ts = static_cast<float*>(workspace);
  float* b_ts = a_ts + d.a_floats;
  int kb0 = 0, njobs = 0, gx = 1;
  Ts32Jobs jobs{};
  for (int s = 0; s < g->nseg; ++s) {
    const l2hmc_gemm_seg& sg = g->seg[s];
    const int nkb = (sg.K + 31) / 32;
    if (nkb == 0) continue;
    if (a_ts_in == nullptr) {
      jobs.j[njobs++] = Ts32Job{sg.A, sg.a_rs, sg.a_cs, g->M, sg.K, a_ts, d.rows_p / 8, kb0, nkb};
      gx = std::max(gx, ew_grid(static_cast<size_t>(nkb) * (d.rows_p / 8) * 64));
    }
    // B operand rows are the output columns: element (n, k) = B[k * b_rs + n * b_cs]
    jobs.j[njobs++] = Ts32Job{sg.B, sg.b_cs, sg.b_rs, g->N, sg.K, b_ts, d.n_p / 8, kb0, nkb};
    gx = std::max(gx, ew_grid(static_cast<size_t>(nkb) * (d.n_p / 8) * 64));
    kb0 += nkb;
  }
  int rc = L2HMC_OK;
  if (njobs > 0) {
    launch_pdl(ts32_from_strided_kernel, dim3(gx, njobs), dim3(256), 0, st, jobs);
    rc = check_launch("ts32_from_strided_kernel");
  }
  if (rc) return rc;
  if (d.kb == 0) {
    set_error("l2hmc_tc_gemm: empty K");
    return L2HMC_ERR_INVALID;
  }
  TcParams p{};
  p.A = a_ts_in ? a_ts_in : a_ts, p.a_rg = d.rows_p / 8, p.B = b_ts, p.b_rg = d.n_p / 8, p.KB = d.kb;
  p.out_ts = out_ts, p.out_rg = d.rows_p / 8, p.out_kb = (g->N + 31) / 32;
  p.m_tiles = d.rows_p / 128;
  p.rows = g->M, p.cols = g->N;
  p.C = g->C, p.ldc = g->ldc, p.accumulate = g->accumulate, p.gate = g->relu_gate, p.ldg = g->ldg;
  p.cscale = g->col_scale, p.cinv = g->col_scale_invert, p.row_set = g->row_set;
  p.cscale_alt = g->col_scale_alt, p.cinv_alt = g->col_scale_invert_alt;
  p.n_tiles = d.n_tiles;
  p.ksplit = d.ksplit, p.kb_per = d.kb_per, p.partial = b_ts + d.b_floats;
  rc = d.narrow ? launch_tc<128, kTcPlain>(p, st) : launch_tc<256, kTcPlain>(p, st);
  if (rc || d.ksplit <= 1) return rc;
  return launch_splitk_reduce(p.partial, d.ksplit, g, st);
}

// Dispatch of the training GEMMs (north star: tensor cores only where the product is a dense contraction, FFMA
// otherwise): tcgen05 3xTF32 when every dimension fills MMA tiles and the product is large enough to amortise the two
// operand conversions; l2hmc_sgemm (ksplit as given) for the small-batch, M = 1 / M = 2 and ragged cases.
// L2HMC_TRAIN_GEMM=ffma forces the FFMA kernel (diagnostics).
static bool gemm_wants_tc(const l2hmc_gemm_args* g) {
  static const bool force_ffma = getenv("L2HMC_TRAIN_GEMM") && std::string(getenv("L2HMC_TRAIN_GEMM")) == "ffma";
  if (force_ffma || g->M < 128 || g->N < 64) return false;
  long long k = 0;
  for (int s = 0; s < g->nseg; ++s) k += g->seg[s].K;
  if (k < 128) return false;
  static const double min_mnk = getenv("L2HMC_TC_MIN_MNK") ? atof(getenv("L2HMC_TC_MIN_MNK")) : 2.5e8;   // diagnostics
  return static_cast<double>(g->M) * g->N * static_cast<double>(k) >= min_mnk;
}

// this product runs on tcgen05 under l2hmc_gemm's dispatch and is not split over K: it can be a link of tc_gemm_chain
bool l2hmc::tc_gemm_chainable(const l2hmc_gemm_args* g) {
  return g && g->M > 0 && g->N > 0 && g->nseg >= 1 && g->nseg <= 3 && gemm_wants_tc(g) && tc_gemm_dims(g).ksplit == 1;
}

extern "C" size_t l2hmc_gemm_workspace_bytes(const l2hmc_gemm_args* g, int ksplit) {
  if (!g || g->nseg < 1 || g->nseg > 3 || g->M <= 0 || g->N <= 0) return 0;
  return gemm_wants_tc(g) ? l2hmc_tc_gemm_workspace_bytes(g) : l2hmc_sgemm_workspace_bytes(g, ksplit);
}

extern "C" int l2hmc_gemm(const l2hmc_gemm_args* g, int ksplit, void* workspace, size_t workspace_bytes, void* stream) {
  L2HMC_REQUIRE(g && g->nseg >= 1 && g->nseg <= 3, "l2hmc_gemm: bad arguments");
  if (g->M > 0 && g->N > 0 && gemm_wants_tc(g)) return l2hmc_tc_gemm(g, workspace, workspace_bytes, stream);
  return l2hmc_sgemm(g, ksplit, workspace, workspace_bytes, stream);
}

// Diagnostic: C = relu(A . W) through the tensor-core layer kernel and the TS32 converters
// (A [M,K] and W [K,Nc] row-major fp32 in, C [M,Nc] row-major out; scratch >= l2hmc_tc_selftest_bytes).
extern "C" size_t l2hmc_tc_selftest_bytes(int M, int Nc, int K) {
  const size_t rows_p = (M + 127) / 128 * 128, n_p = (Nc + 255) / 256 * 256, kb = (K + 31) / 32, kbn = (Nc + 31) / 32;
  return sizeof(float) * (kb * rows_p * 64 + kb * n_p * 64 + kbn * rows_p * 64 + n_p);
}

extern "C" int l2hmc_tc_selftest(const float* A, const float* W, int M, int Nc, int K, int bn, float* C, void* scratch,
                                 void* stream) {
  L2HMC_REQUIRE(A && W && C && scratch, "l2hmc_tc_selftest: null pointer");
  L2HMC_REQUIRE(M > 0 && Nc > 0 && K > 0 && (bn == 128 || bn == 256), "l2hmc_tc_selftest: bad sizes");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rows_p = (M + 127) / 128 * 128, n_p = (Nc + 255) / 256 * 256, kb = (K + 31) / 32, kbn = (Nc + 31) / 32;
  float* a_ts = static_cast<float*>(scratch);
  float* w_ts = a_ts + static_cast<size_t>(kb) * rows_p * 64;
  float* c_ts = w_ts + static_cast<size_t>(kb) * n_p * 64;
  float* bias = c_ts + static_cast<size_t>(kbn) * rows_p * 64;
  cudaMemsetAsync(w_ts, 0, sizeof(float) * static_cast<size_t>(kb) * n_p * 64, st);
  cudaMemsetAsync(bias, 0, sizeof(float) * n_p, st);
  ts32_from_rowmajor_kernel<<<ew_grid(static_cast<size_t>(kb) * (rows_p / 8) * 64), 256, 0, st>>>(
      A, K, M, K, nullptr, 0, nullptr, 0, nullptr, a_ts, rows_p / 8, 0, kb);
  ts32_pack_weight_kernel<<<ew_grid(static_cast<size_t>(kb) * 32 * Nc), 256, 0, st>>>(W, K, Nc, w_ts, n_p / 8, 0, 0, 0);
  TcParams p{};
  p.A = a_ts, p.a_rg = rows_p / 8, p.B = w_ts, p.b_rg = n_p / 8, p.KB = kb;
  p.m_tiles = rows_p / 128, p.n_tiles = (Nc + bn - 1) / bn;
  p.rows = M, p.cols = Nc, p.bias = bias, p.out_ts = c_ts, p.out_rg = rows_p / 8, p.out_kb = kbn;
  int rc = bn == 128 ? launch_tc<128, kTcLayer>(p, st) : launch_tc<256, kTcLayer>(p, st);
  if (rc) return rc;
  ts32_to_rowmajor_kernel<<<ew_grid(static_cast<size_t>(M) * Nc), 256, 0, st>>>(c_ts, rows_p / 8, M, Nc, C);
  return check_launch("l2hmc_tc_selftest", 3);
}
