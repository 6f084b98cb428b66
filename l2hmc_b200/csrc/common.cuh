// Shared device/host helpers for the l2hmc_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/l2hmc_b200.h"

namespace l2hmc {

// fp32 constants exactly as TF folds the reference's python scalars into fp32 tensors.
constexpr float kPi = 3.14159265358979323846f;
constexpr float kTwoPi = 6.28318530717958647692f;

void set_error(const char* fmt, ...);
int check_launch(const char* what, int launches = 1);   // also adds `launches` to l2hmc_launch_count()
// C (=|+=) epilogue(sum_z partial[z][M][N]), z in fixed order: the second stage of every split-K product (train.cu)
int launch_splitk_reduce(const float* partial, int ksplit, const l2hmc_gemm_args* g, cudaStream_t stream);
// Several column sums of matrices with the same row count in two launches (train.cu); every sum is formed exactly as
// l2hmc_col_sum forms it.  out[1], out[2] (nullable) receive copies of the sums.
struct ColSumJob {
  const float* B;
  long long ld;
  int cols;
  float* out[3];
};
constexpr int kMaxColSumJobs = 8;
size_t col_sum_multi_workspace_bytes(int rows, int max_cols, int njobs);
int col_sum_multi(const ColSumJob* jobs, int njobs, int rows, void* workspace, size_t workspace_bytes, cudaStream_t stream);
// l2hmc_train_update_fwd with the step size optionally read from device memory and with extra outputs: the network
// inputs of the calls that consume the new state, written by the kernel that produces it (train.cu; used by
// train_step.cu instead of a copy / mask-product launch per call)
struct UpdExtra {
  float* copy[2];          // nullable: further copies of `out` (the in0 slots of the tape)
  float* masked;           // nullable: nmask (*) out with the per-chain mask set (the in1 slot of the next position call)
  const float* nmask;
  const float* nmask_alt;
  int ninv, ninv_alt;
};
int train_update_fwd(int position, const float* x, const float* v, const float* g, const float* S, const float* Tr,
                     const float* Q, const float* mask, int invert, const float* mask_alt, int invert_alt,
                     const uint8_t* row_set, float eps, const float* eps_dev, int dir, int dir_alt, int N, int D, float* out,
                     float* logdet, int32_t* nonfinite_flag, const UpdExtra* extra, void* stream);

// ---- programmatic dependent launch (PDL).  A kernel that starts with pdl_wait() may be launched with launch_pdl():
// its CTAs become resident (launch latency, barrier / TMEM set-up before the wait) while the kernel ahead of it in the
// stream is still running, once every CTA of THAT kernel has called pdl_trigger() or exited; pdl_wait() returns when
// the kernels ahead have completed and their writes are visible.  Both instructions are no-ops in a kernel launched
// without the attribute, and a kernel that never triggers simply releases its dependents at completion, so the two
// halves can be adopted kernel by kernel.  Every thread calls pdl_wait() before its first global access (so a grid
// never completes before the grids ahead of it).  L2HMC_PDL=0 launches everything without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr, cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

#define L2HMC_REQUIRE(cond, ...)          \
  do {                                    \
    if (!(cond)) {                        \
      ::l2hmc::set_error(__VA_ARGS__);    \
      return L2HMC_ERR_INVALID;           \
    }                                     \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ------------------------------------------------------------------ warp helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <int W>
__device__ __forceinline__ float group_sum(float v) {  // sum over aligned groups of W lanes
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------ mbarrier / bulk-copy PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a mis-armed barrier traps (launch error) instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 24)) __trap();
  }
}
// Same, for a warp that has nothing else to do (the copy warp): sleep between polls so the
// spin does not take issue slots from the compute warps of the SM.
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(256);
    if (++spins > (1u << 22)) __trap();
  }
}
// 1-D bulk async copy global -> shared (TMA engine, no tensor map); completes on `bar`.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// 1-D bulk async copy shared -> global.
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy smem writes -> visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ packed fp32x2 (Blackwell FFMA2/FADD2/FMUL2)
// Two IEEE fp32 lanes per instruction, each rounded exactly like the scalar op: the issue-bound
// fused kernels do their element-wise math on pairs.  NOTE ptxas contracts mul.rn.f32x2 followed
// by add.rn.f32x2 into one FFMA2 (even with explicit .rn); where the reference's separately
// rounded `a + b*c` is wanted, use madd2_unfused(), whose add is an FFMA2 by a RUNTIME 1.0.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float lo2(f32x2 v) { return __uint_as_float(static_cast<unsigned>(v)); }
__device__ __forceinline__ float hi2(f32x2 v) { return __uint_as_float(static_cast<unsigned>(v >> 32)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// a + b*c with the product and the sum rounded separately (`one` must be a runtime (1,1)).
__device__ __forceinline__ f32x2 madd2_unfused(f32x2 a, f32x2 b, f32x2 c, f32x2 one) {
  return fma2(a, one, mul2(b, c));
}

// ------------------------------------------------------------------ reference fp32 idioms
// p - 2pi*floor((p + pi)/2pi) with every op rounded separately (lattice/lattice.py:47-49).
__device__ __forceinline__ float project_angle(float p) {
  float q = floorf(__fdiv_rn(__fadd_rn(p, kPi), kTwoPi));
  return __fsub_rn(p, __fmul_rn(kTwoPi, q));
}
// numpy's float32 np.mod(x, 2pi): fmod, then shift negatives up by the divisor.
__device__ __forceinline__ float wrap_2pi(float x) {
  const float ax = fabsf(x);
  float r;
  if (__builtin_expect(ax < 65536.f, 1)) {
    // exact fmod(|x|, c): for q = floor(|x|/c) the value |x| - q*c is a multiple of 2^-21 below 8,
    // hence representable, and one FMA delivers it exactly; q from the reciprocal can be off by one.
    float q = floorf(__fmul_rn(ax, 0.15915494309189535f));
    r = fmaf(-q, kTwoPi, ax);
    if (r < 0.f) r = fmaf(-(q - 1.f), kTwoPi, ax);
    else if (r >= kTwoPi) r = fmaf(-(q + 1.f), kTwoPi, ax);
  } else {
    r = fmodf(ax, kTwoPi);
  }
  if (x < 0.f && r != 0.f) r = __fadd_rn(-r, kTwoPi);   // numpy: negative remainder + divisor
  return r;
}

// ------------------------------------------------------------------ Philox4x32-10 (counter-based RNG)
struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  __device__ __host__ static inline void run(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                             uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      uint64_t p0 = static_cast<uint64_t>(M0) * c0;
      uint64_t p1 = static_cast<uint64_t>(M1) * c2;
      uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k0;
      uint32_t n1 = static_cast<uint32_t>(p1);
      uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k1;
      uint32_t n3 = static_cast<uint32_t>(p0);
      c0 = n0, c1 = n1, c2 = n2, c3 = n3;
      k0 += W0, k1 += W1;
    }
    out[0] = c0, out[1] = c1, out[2] = c2, out[3] = c3;
  }
};
// Counter layout (documented in DESIGN.md): c0 = index inside the chain (link quad, or
// 0xFFFFFFFF for the per-chain uniforms), c1 = global chain index (low 32 bits),
// c2 = MC step (low 32), c3 = chain index high 16 | MC step high 16.  key = seed.
__device__ __forceinline__ void philox_chain(uint64_t seed, uint64_t chain, uint64_t mc_step, uint32_t idx,
                                             uint32_t out[4]) {
  Philox::run(idx, static_cast<uint32_t>(chain), static_cast<uint32_t>(mc_step),
              (static_cast<uint32_t>(chain >> 32) << 16) | (static_cast<uint32_t>(mc_step >> 32) & 0xffffu),
              static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32), out);
}
__device__ __forceinline__ float u01(uint32_t r) { return static_cast<float>(r >> 8) * 5.9604644775390625e-08f; }
// Box-Muller: two uniforms -> two standard normals.
__device__ __forceinline__ void box_muller(uint32_t r0, uint32_t r1, float& z0, float& z1) {
  float u1 = static_cast<float>((r0 >> 8) + 1u) * 5.9604644775390625e-08f;  // (0,1]
  float u2 = u01(r1);
  float rad = sqrtf(-2.f * __logf(u1));        // MUFU.LG2 / MUFU.SIN / MUFU.COS: plenty for a generator
  float s, c;
  __sincosf(kTwoPi * u2 - kPi, &s, &c);        // argument in [-pi, pi): the intrinsics' accurate range
  z0 = rad * c;
  z1 = rad * s;
}
// The four normals of link quad `quad` (links 4*quad .. 4*quad+3 of the chain).
__device__ __forceinline__ void philox_normal4(uint64_t seed, uint64_t chain, uint64_t mc_step, uint32_t quad,
                                               float z[4]) {
  uint32_t r[4];
  philox_chain(seed, chain, mc_step, quad, r);
  box_muller(r[0], r[1], z[0], z[1]);
  box_muller(r[2], r[3], z[2], z[3]);
}
__device__ __forceinline__ void philox_uniform2(uint64_t seed, uint64_t chain, uint64_t mc_step, float& u_dir,
                                                float& u_acc) {
  uint32_t r[4];
  philox_chain(seed, chain, mc_step, 0xffffffffu, r);
  u_dir = u01(r[0]);
  u_acc = u01(r[1]);
}

}  // namespace l2hmc
