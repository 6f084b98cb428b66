// SURVEY section 8(f) rank 1 (forward part): the training loss VALUE of the reference, per chain and averaged.
//
// Replaces (reference l2hmc/gauge_model.py):
//   :94-108    project_angle_approx (Fourier series of the sawtooth, harmonics 1..4)
//   :631-657   metric functions l1 / l2 / cos / cos2 / cos_diff
//   :699-725   _calc_top_charges(fft) / _calc_top_charges_diff
//   :762-801   the loss terms of _calc_loss (std loss on the metric + charge loss, x-chain and auxiliary z-chain;
//              the z-terms compare z with the x-chain's proposal, as the reference does)
//
// One warp owns one chain: x, x_prop and z are staged once through shared memory (their plaquette angles need
// neighbours), the metric sums are taken from the values while they pass through registers, x_out re-uses the first
// buffer for the integer charge difference.  Per-chain sums are lane-strided partials closed by a shuffle tree;
// the batch mean is a second, single-block kernel with a fixed tree (deterministic).
// HBM traffic: 16 B/link read (x, x_prop, z, x_out), O(N) written.
#include "common.cuh"
#include "fast_sin.cuh"

namespace l2hmc {
namespace {

struct LossParams {
  const float* x;
  const float* x_prop;
  const float* z;
  const float* x_out;
  const float* px;
  const float* pz;
  int N, T, X;
  int metric;
  float loss_scale, aux_weight, std_weight, charge_weight;
  float* per_chain;
  int32_t* x_dq;
  float* q_fft;   // with `only_fft`: sum proj_approx(P) / 2pi of x
};

__device__ __forceinline__ float metric_elem(int metric, float a, float b) {
  switch (metric) {
    case L2HMC_METRIC_L1: return fabsf(__fsub_rn(a, b));
    case L2HMC_METRIC_L2: { const float d = __fsub_rn(a, b); return __fmul_rn(d, d); }
    case L2HMC_METRIC_COS: return fabsf(__fsub_rn(fast_cos(a), fast_cos(b)));
    case L2HMC_METRIC_COS2: { const float d = __fsub_rn(fast_cos(a), fast_cos(b)); return __fmul_rn(d, d); }
    default: return __fsub_rn(1.f, fast_cos(__fsub_rn(a, b)));   // cos_diff
  }
}

// sum_{n=1}^{4} (-2/n)(-1)^n sin(n p), accumulated in the reference's order
__device__ __forceinline__ float project_angle_approx(float p) {
  float y = __fmul_rn(2.f, fast_sin(p));
  y = __fadd_rn(y, __fmul_rn(-1.f, fast_sin(__fmul_rn(2.f, p))));
  y = __fadd_rn(y, __fmul_rn(0.6666666666666666f, fast_sin(__fmul_rn(3.f, p))));
  y = __fadd_rn(y, __fmul_rn(-0.5f, fast_sin(__fmul_rn(4.f, p))));
  return y;
}

template <bool FFT>
__device__ __forceinline__ float charge_sum(const float* s, int T, int X, int lane) {
  const int SITES = T * X;
  float q = 0.f;
  for (int i = lane; i < SITES; i += 32) {
    const int t = i / X, xx = i - t * X;
    const int sr = t * X + (xx + 1 == X ? 0 : xx + 1);
    const int sd = (t + 1 == T ? 0 : t + 1) * X + xx;
    const float P = __fadd_rn(__fsub_rn(__fsub_rn(s[2 * i], s[2 * i + 1]), s[2 * sr]), s[2 * sd + 1]);
    q = __fadd_rn(q, FFT ? project_angle_approx(P) : project_angle(P));
  }
  return __fdiv_rn(warp_sum(q), kTwoPi);
}

__global__ void __launch_bounds__(128) loss_kernel(const LossParams p, int warps_per_block, int only_fft) {
  extern __shared__ __align__(16) float s_all[];
  const int T = p.T, X = p.X, D = 2 * T * X;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Dp = (D + 1) & ~1;
  float* s0 = s_all + static_cast<size_t>(warp) * 3 * Dp;
  float* s1 = s0 + Dp;
  float* s2 = s1 + Dp;
  for (int n = blockIdx.x * warps_per_block + warp; n < p.N; n += gridDim.x * warps_per_block) {
    const size_t base = static_cast<size_t>(n) * D;
    if (only_fft) {
      for (int i = lane; i < D; i += 32) s0[i] = p.x[base + i];
      __syncwarp();
      const float q = charge_sum<true>(s0, T, X, lane);
      if (lane == 0) p.q_fft[n] = q;
      __syncwarp();
      continue;
    }
    float xs = 0.f, zs = 0.f;
    for (int i = lane; i < D; i += 32) {
      const float a = p.x[base + i], b = p.x_prop[base + i], c = p.z[base + i];
      s0[i] = a, s1[i] = b, s2[i] = c;
      xs = __fadd_rn(xs, metric_elem(p.metric, a, b));
      zs = __fadd_rn(zs, metric_elem(p.metric, c, b));
    }
    __syncwarp();
    xs = warp_sum(xs);
    zs = warp_sum(zs);
    const float qx_fft = charge_sum<true>(s0, T, X, lane);
    const float qp_fft = charge_sum<true>(s1, T, X, lane);
    const float qz_fft = charge_sum<true>(s2, T, X, lane);
    const float qx = charge_sum<false>(s0, T, X, lane);
    __syncwarp();
    for (int i = lane; i < D; i += 32) s0[i] = p.x_out[base + i];
    __syncwarp();
    const float qo = charge_sum<false>(s0, T, X, lane);
    if (lane == 0) {
      const float eps = 1e-3f;
      const float px = p.px[n], pz = p.pz[n];
      const float x_std = __fadd_rn(__fmul_rn(xs, px), eps);
      const float z_std = __fmul_rn(p.aux_weight, __fadd_rn(__fmul_rn(zs, pz), eps));
      const float inv = __fadd_rn(__fdiv_rn(1.f, x_std), __fdiv_rn(1.f, z_std));
      const float std_loss = __fmul_rn(
          __fsub_rn(__fmul_rn(p.loss_scale, inv), __fdiv_rn(__fadd_rn(x_std, z_std), p.loss_scale)), p.std_weight);
      const float xq = __fadd_rn(__fmul_rn(px, fabsf(__fsub_rn(qx_fft, qp_fft))), eps);
      const float zq = __fmul_rn(p.aux_weight, __fadd_rn(__fmul_rn(pz, fabsf(__fsub_rn(qz_fft, qp_fft))), eps));
      p.per_chain[n] = __fadd_rn(std_loss, __fmul_rn(p.charge_weight, __fadd_rn(xq, zq)));
      if (p.x_dq) p.x_dq[n] = static_cast<int32_t>(fabsf(__fsub_rn(qx, qo)));   // tf.cast(|dQ|, int32)
    }
    __syncwarp();
  }
}

// out[0] = mean(v[0..N)) with a fixed summation tree (one block)
__global__ void __launch_bounds__(1024) mean_kernel(const float* __restrict__ v, int N, float* out) {
  __shared__ float part[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < N; i += 1024) s = __fadd_rn(s, v[i]);
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = warp_sum(part[threadIdx.x]);
    if (threadIdx.x == 0) out[0] = __fdiv_rn(s, static_cast<float>(N));
  }
}

int launch_loss(const LossParams& p, int only_fft, cudaStream_t st) {
  const size_t per_warp = sizeof(float) * 3 * ((2u * p.T * p.X + 1u) & ~1u);
  int dev = 0, max_smem = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int warps = 4;
  while (warps > 1 && per_warp * warps > 100 * 1024) warps >>= 1;
  const size_t smem = per_warp * warps;
  if (smem > static_cast<size_t>(max_smem)) {
    set_error("l2hmc_u1_loss: a %dx%d chain (%zu B) does not fit shared memory (%d B)", p.T, p.X, smem, max_smem);
    return L2HMC_ERR_UNSUPPORTED;
  }
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(loss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(loss_kernel): %s", cudaGetErrorString(e));
      return L2HMC_ERR_CUDA;
    }
  }
  const int grid = max(1, min((p.N + warps - 1) / warps, sms * 8));
  loss_kernel<<<grid, warps * 32, smem, st>>>(p, warps, only_fft);
  return check_launch("loss_kernel");
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_u1_loss(const l2hmc_loss_params* prm, const float* x, const float* x_prop, const float* z,
                             const float* x_out, const float* px, const float* pz, int N, int T, int X,
                             float* per_chain, int32_t* x_dq, float* loss, void* stream) {
  L2HMC_REQUIRE(prm, "l2hmc_u1_loss: null parameters");
  L2HMC_REQUIRE(N > 0 && T > 0 && X > 0, "l2hmc_u1_loss: bad sizes N=%d T=%d X=%d", N, T, X);
  L2HMC_REQUIRE(x && x_prop && z && x_out && px && pz && per_chain, "l2hmc_u1_loss: null pointer");
  L2HMC_REQUIRE(prm->metric >= L2HMC_METRIC_L1 && prm->metric <= L2HMC_METRIC_COS_DIFF, "l2hmc_u1_loss: metric %d",
                prm->metric);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  LossParams p{x, x_prop, z, x_out, px, pz, N, T, X, prm->metric, prm->loss_scale, prm->aux_weight, prm->std_weight,
               prm->charge_weight, per_chain, x_dq, nullptr};
  int rc = launch_loss(p, 0, st);
  if (rc || !loss) return rc;
  mean_kernel<<<1, 1024, 0, st>>>(per_chain, N, loss);
  return check_launch("mean_kernel");
}

extern "C" int l2hmc_u1_top_charges_fft(const float* x, int N, int T, int X, float* q, void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(x && q, "l2hmc_u1_top_charges_fft: null pointer");
  L2HMC_REQUIRE(N > 0 && T > 0 && X > 0, "l2hmc_u1_top_charges_fft: bad sizes N=%d T=%d X=%d", N, T, X);
  LossParams p{};
  p.x = x, p.N = N, p.T = T, p.X = X, p.q_fft = q;
  return launch_loss(p, 1, static_cast<cudaStream_t>(stream));
}
