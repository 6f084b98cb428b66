// run-loop statistics as streaming device-side accumulators.
//
// Replaces (reference l2hmc/...):
//   gauge_model.py:1473-1533   calc_observables_stats: per-chain mean and scipy.stats.sem over the MC steps of
//                              action / plaquette / charge / charge^2 after a thermalisation cut, and the
//                              relative frequency of every integer charge
//   utils/trajectories.py:63-95  calc_tunneling_rate: mode switches between consecutive steps
//
// The reference keeps every step's (N,) arrays on the host and reduces at the end; here each MC step folds its
// observables into fp64 running sums (sum, sum of squares) per chain and an integer histogram, so nothing but the
// final (4, N) means / errors ever leaves the GPU.  Integer atomics only (order-independent, deterministic).
// Quirks kept: charges are truncated toward zero like `np.array(..., dtype=int)` (:1496), and charge^2 is
// accumulated over ALL steps while the other three skip the thermalisation steps (:1498 precedes the cut :1503).
#include "common.cuh"

namespace l2hmc {
namespace {

__global__ void __launch_bounds__(256) stats_update_kernel(const l2hmc_run_stats s, const float* __restrict__ action,
                                                            const float* __restrict__ plaq,
                                                            const float* __restrict__ charge, int counted) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= s.N) return;
  const long long q = static_cast<long long>(charge[n]);           // truncation toward zero
  const double q2 = static_cast<double>(q * q);
  s.sum[3 * static_cast<size_t>(s.N) + n] += q2;
  s.sumsq[3 * static_cast<size_t>(s.N) + n] += q2 * q2;
  if (!counted) return;
  const double a = static_cast<double>(action[n]), p = static_cast<double>(plaq[n]), qd = static_cast<double>(q);
  s.sum[n] += a, s.sumsq[n] += a * a;
  s.sum[static_cast<size_t>(s.N) + n] += p, s.sumsq[static_cast<size_t>(s.N) + n] += p * p;
  s.sum[2 * static_cast<size_t>(s.N) + n] += qd, s.sumsq[2 * static_cast<size_t>(s.N) + n] += qd * qd;
  if (s.hist) {
    const long long bin = q + s.qmax;
    const int slot = (bin < 0 || bin > 2LL * s.qmax) ? 2 * s.qmax + 1 : static_cast<int>(bin);   // last slot: overflow
    atomicAdd(&s.hist[slot], 1ULL);
  }
}

// mean = sum / n, err = sqrt(var(ddof = 1) / n); rows 0-2 over the counted steps, row 3 over all steps
__global__ void __launch_bounds__(256) stats_finalize_kernel(const l2hmc_run_stats s, long long n_counted,
                                                              long long n_total, double* mean, double* err) {
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= 4 * static_cast<size_t>(s.N)) return;
  const double n = static_cast<double>(i >= 3 * static_cast<size_t>(s.N) ? n_total : n_counted);
  const double m = s.sum[i] / n;
  mean[i] = m;
  const double var = n > 1.0 ? fmax((s.sumsq[i] - s.sum[i] * m) / (n - 1.0), 0.0) : nan("");
  err[i] = sqrt(var / n);
}

__global__ void __launch_bounds__(256) mode_switch_kernel(const int32_t* __restrict__ mode, int32_t* prev,
                                                           uint32_t* switches, int have_prev, int N) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int32_t m = mode[n];
  if (have_prev && prev[n] != m) switches[n] += 1u;
  prev[n] = m;
}

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" int l2hmc_stats_update(const l2hmc_run_stats* s, const float* action, const float* plaq,
                                  const float* charge, int counted, void* stream) {
  L2HMC_REQUIRE(s && s->sum && s->sumsq, "l2hmc_stats_update: null accumulator");
  if (s->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(s->N > 0 && s->qmax >= 0, "l2hmc_stats_update: bad sizes N=%d qmax=%d", s->N, s->qmax);
  L2HMC_REQUIRE(charge && (!counted || (action && plaq)), "l2hmc_stats_update: null observable");
  stats_update_kernel<<<(s->N + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(*s, action, plaq, charge,
                                                                                         counted);
  return check_launch("stats_update_kernel");
}

extern "C" int l2hmc_stats_finalize(const l2hmc_run_stats* s, long long n_counted, long long n_total, double* mean,
                                    double* err, void* stream) {
  L2HMC_REQUIRE(s && s->sum && s->sumsq && mean && err, "l2hmc_stats_finalize: null pointer");
  if (s->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(n_counted > 0 && n_total >= n_counted, "l2hmc_stats_finalize: bad step counts %lld / %lld", n_counted,
                n_total);
  const size_t total = 4 * static_cast<size_t>(s->N);
  stats_finalize_kernel<<<static_cast<int>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      *s, n_counted, n_total, mean, err);
  return check_launch("stats_finalize_kernel");
}

extern "C" int l2hmc_mode_switches(const int32_t* mode, int32_t* prev, uint32_t* switches, int have_prev, int N,
                                   void* stream) {
  if (N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(mode && prev && switches && N > 0, "l2hmc_mode_switches: bad argument");
  mode_switch_kernel<<<(N + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(mode, prev, switches, have_prev, N);
  return check_launch("mode_switch_kernel");
}
