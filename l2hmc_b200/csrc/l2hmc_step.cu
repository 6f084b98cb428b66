// The whole L2HMC Monte-Carlo step (with S/T/Q networks) enqueued by ONE host call.
//
// Replaces the host-side control flow of (reference l2hmc/dynamics/gauge_dynamics.py)
//   :195-259   apply_transition (direction draw, the two transition kernels, blends, accept / reject)
//   :261-313   transition_kernel (momentum draw, tf.while_loop over the leapfrog steps, accept probability)
//   :412-483   _forward_lf / _backward_lf (v-half, x(m), x(mbar), v-half)
// and gauge_model.py:1180,1388 (np.mod 2 pi) / :683-716 (observables of the new state) when asked to.
// It launches only kernels of this library, in stream order, on caller-owned memory: about 20 launches per
// leapfrog step issued back to back from C++ (a Python loop pays ~20 us of interpreter + ctypes time per
// launch, which is the whole cost of a small-batch step).  Every chain runs only the direction it drew; both
// kinds share each launch (l2hmc_net_call.row_set).  Results are identical to the kernel-by-kernel composition
// (tests/test_gpu_l2hmc.py::test_native_mc_step_equals_python_composition).
#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace l2hmc {
namespace {

// row_set[n] = 1 for chains that integrate backward (u_dir <= 0.5: the reference's forward mask is u > 0.5)
__global__ void __launch_bounds__(256) row_set_kernel(const float* __restrict__ u_dir, int N, uint8_t* row_set) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < N) row_set[n] = u_dir[n] > 0.5f ? 0 : 1;
}

struct Carve {
  char* p;
  size_t left;
  template <typename T>
  T* take(size_t count) {
    const size_t bytes = (count * sizeof(T) + 255) & ~static_cast<size_t>(255);
    if (bytes > left) return nullptr;
    T* r = reinterpret_cast<T*>(p);
    p += bytes, left -= bytes;
    return r;
  }
};

size_t align256(size_t b) { return (b + 255) & ~static_cast<size_t>(255); }

}  // namespace
}  // namespace l2hmc

using namespace l2hmc;

extern "C" size_t l2hmc_u1_l2hmc_mc_workspace_bytes(const l2hmc_l2hmc_mc_args* a) {
  if (!a || !a->xnet || !a->vnet || a->N <= 0 || a->T <= 0 || a->X <= 0) return 0;
  const size_t N = a->N, D = 2 * static_cast<size_t>(a->T) * a->X;
  size_t total = align256(N * D * 4)      // force
                 + 3 * align256(N * 4)    // h_old, h_new, sumlogdet
                 + 2 * align256(N * 4)    // u_dir, u_acc (Philox mode)
                 + align256(N);           // row_set
  const size_t nw = std::max(l2hmc_net_workspace_bytes(a->xnet->K, a->xnet->H, a->xnet->D, a->N),
                             l2hmc_net_workspace_bytes(a->vnet->K, a->vnet->H, a->vnet->D, a->N));
  total += align256(nw);
  if (a->xconv_in0) total += 2 * align256(N * static_cast<size_t>(a->xnet->K) * 4);   // conv features of both streams
  return total + 256;
}

extern "C" int l2hmc_u1_l2hmc_mc_step(const l2hmc_l2hmc_mc_args* a, void* stream) {
  L2HMC_REQUIRE(a, "l2hmc_u1_l2hmc_mc_step: null arguments");
  if (a->N == 0) return L2HMC_OK;
  L2HMC_REQUIRE(a->N > 0 && a->T > 0 && a->X > 0 && a->M > 0, "l2hmc_u1_l2hmc_mc_step: bad sizes N=%d T=%d X=%d M=%d",
                a->N, a->T, a->X, a->M);
  L2HMC_REQUIRE(a->xnet && a->vnet && a->masks && a->ts && a->x, "l2hmc_u1_l2hmc_mc_step: null input");
  L2HMC_REQUIRE(a->x_post && a->v_post && a->accept_prob, "l2hmc_u1_l2hmc_mc_step: x_post, v_post, accept_prob required");
  const bool conv = a->xconv_in0 != nullptr;
  L2HMC_REQUIRE(conv == (a->xconv_in1 != nullptr) && conv == (a->vconv_in0 != nullptr) &&
                    conv == (a->vconv_in1 != nullptr),
                "l2hmc_u1_l2hmc_mc_step: the four conv weight sets must be given together");
  const bool inject = a->v_f != nullptr;
  L2HMC_REQUIRE(!inject || (a->u_dir && a->u_acc), "l2hmc_u1_l2hmc_mc_step: injected momenta need u_dir and u_acc");
  const int N = a->N, T = a->T, X = a->X, M = a->M, D = 2 * T * X;
  L2HMC_REQUIRE(a->xnet->D == D && a->vnet->D == D, "l2hmc_u1_l2hmc_mc_step: network output width != 2TX");
  L2HMC_REQUIRE(a->workspace && a->workspace_bytes >= l2hmc_u1_l2hmc_mc_workspace_bytes(a),
                "l2hmc_u1_l2hmc_mc_step: workspace too small (%zu < %zu bytes)", a->workspace_bytes,
                l2hmc_u1_l2hmc_mc_workspace_bytes(a));
  L2HMC_REQUIRE((reinterpret_cast<uintptr_t>(a->workspace) & 255u) == 0, "l2hmc_u1_l2hmc_mc_step: workspace must be "
                "256-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  Carve ws{static_cast<char*>(a->workspace), a->workspace_bytes};
  float* grad = ws.take<float>(static_cast<size_t>(N) * D);
  float* h_old = ws.take<float>(N);
  float* h_new = ws.take<float>(N);
  float* sld = ws.take<float>(N);
  float* u_dir_ws = ws.take<float>(N);
  float* u_acc_ws = ws.take<float>(N);
  uint8_t* row_set = ws.take<uint8_t>(N);
  const size_t nw = std::max(l2hmc_net_workspace_bytes(a->xnet->K, a->xnet->H, a->xnet->D, N),
                             l2hmc_net_workspace_bytes(a->vnet->K, a->vnet->H, a->vnet->D, N));
  char* net_ws = ws.take<char>(nw);
  float* feat0 = conv ? ws.take<float>(static_cast<size_t>(N) * a->xnet->K) : nullptr;
  float* feat1 = conv ? ws.take<float>(static_cast<size_t>(N) * a->xnet->K) : nullptr;
  L2HMC_REQUIRE(net_ws && (!conv || feat1), "l2hmc_u1_l2hmc_mc_step: workspace carve failed");

  int rc;
#define L2HMC_TRY(call) \
  do {                  \
    rc = (call);        \
    if (rc) return rc;  \
  } while (0)

  // ---- draws: momenta into v_post, direction / accept uniforms
  const float* u_dir = a->u_dir;
  const float* u_acc = a->u_acc;
  if (!inject) {
    L2HMC_TRY(l2hmc_philox_draws_dev(a->seed, a->mc_step, a->mc_step_dev, a->chain0, N, D, a->v_post, u_dir_ws, u_acc_ws,
                                     st));
    u_dir = u_dir_ws, u_acc = u_acc_ws;
  } else if (a->v_b == nullptr || a->v_b == a->v_f) {
    if (cudaMemcpyAsync(a->v_post, a->v_f, sizeof(float) * static_cast<size_t>(N) * D, cudaMemcpyDeviceToDevice, st) !=
        cudaSuccess) {
      set_error("l2hmc_u1_l2hmc_mc_step: cudaMemcpyAsync(v)");
      return L2HMC_ERR_CUDA;
    }
  } else {   // v0 = v_f where the chain goes forward, v_b otherwise (exact 0/1 blend); h_old doubles as the dummy p
    L2HMC_TRY(l2hmc_mh_select(a->v_f, a->v_f, a->v_b, nullptr, nullptr, u_dir, u_dir, u_dir, nullptr, N, D, 0, a->v_post,
                              nullptr, nullptr, nullptr, nullptr, st));
  }
  row_set_kernel<<<(N + 255) / 256, 256, 0, st>>>(u_dir, N, row_set);
  L2HMC_TRY(check_launch("row_set_kernel"));
  if (cudaMemcpyAsync(a->x_post, a->x, sizeof(float) * static_cast<size_t>(N) * D, cudaMemcpyDeviceToDevice, st) !=
          cudaSuccess ||
      cudaMemsetAsync(sld, 0, sizeof(float) * N, st) != cudaSuccess) {
    set_error("l2hmc_u1_l2hmc_mc_step: cudaMemcpyAsync / cudaMemsetAsync");
    return L2HMC_ERR_CUDA;
  }
  L2HMC_TRY(l2hmc_u1_hamiltonian(a->x, a->v_post, a->beta, N, T, X, h_old, st));

  // ---- trajectory: chains with row_set = 0 take forward step `step`, the others backward step M - 1 - step
  float* x = a->x_post;
  float* v = a->v_post;
  auto momentum = [&](int step) -> int {
    const int ib = M - 1 - step;
    L2HMC_TRY(l2hmc_u1_action_force(x, a->beta, N, T, X, nullptr, grad, st));
    l2hmc_net_call c{};
    const float* in0 = x;
    const float* in1 = grad;
    if (conv) {
      L2HMC_TRY(l2hmc_conv3d_features(a->vconv_in0, x, nullptr, 0, N, T, X, feat0, st));
      L2HMC_TRY(l2hmc_conv3d_features(a->vconv_in1, grad, nullptr, 0, N, T, X, feat1, st));
      in0 = feat0, in1 = feat1;
    }
    c.in0 = in0, c.in1 = in1, c.N = N, c.mode = L2HMC_NET_MOMENTUM;
    c.tau_cos = a->ts[2 * step], c.tau_sin = a->ts[2 * step + 1];
    c.tau_cos_alt = a->ts[2 * ib], c.tau_sin_alt = a->ts[2 * ib + 1];
    c.eps = a->eps, c.dir = 1, c.dir_alt = -1, c.row_set = row_set;
    c.v = v, c.g = grad, c.out = v, c.logdet = sld, c.nonfinite_flag = a->nonfinite_flag;
    c.workspace = net_ws, c.workspace_bytes = nw, c.path = a->net_path;
    return l2hmc_net_run(a->vnet, &c, st);
  };
  auto position = [&](int step, int invert_fwd) -> int {
    const int ib = M - 1 - step;
    const float* mrow = a->masks + static_cast<size_t>(step) * D;
    const float* mrow_alt = a->masks + static_cast<size_t>(ib) * D;
    l2hmc_net_call c{};
    const float* in0 = v;
    const float* in1 = x;
    if (conv) {
      L2HMC_TRY(l2hmc_conv3d_features(a->xconv_in0, v, nullptr, 0, N, T, X, feat0, st));
      L2HMC_TRY(l2hmc_conv3d_features_sets(a->xconv_in1, x, mrow, invert_fwd, mrow_alt, !invert_fwd, row_set, N, T, X,
                                           feat1, st));
      in0 = feat0, in1 = feat1;
    } else {
      c.in1_mask = mrow, c.in1_mask_invert = invert_fwd;
      c.in1_mask_alt = mrow_alt, c.in1_mask_invert_alt = !invert_fwd;
    }
    c.in0 = in0, c.in1 = in1, c.N = N, c.mode = L2HMC_NET_POSITION;
    c.tau_cos = a->ts[2 * step], c.tau_sin = a->ts[2 * step + 1];
    c.tau_cos_alt = a->ts[2 * ib], c.tau_sin_alt = a->ts[2 * ib + 1];
    c.eps = a->eps, c.dir = 1, c.dir_alt = -1, c.row_set = row_set;
    c.x = x, c.v = v, c.mask = mrow, c.mask_invert = invert_fwd, c.mask_alt = mrow_alt, c.mask_invert_alt = !invert_fwd;
    c.out = x, c.logdet = sld, c.nonfinite_flag = a->nonfinite_flag;
    c.workspace = net_ws, c.workspace_bytes = nw, c.path = a->net_path;
    return l2hmc_net_run(a->xnet, &c, st);
  };
  for (int step = 0; step < M; ++step) {
    L2HMC_TRY(momentum(step));
    L2HMC_TRY(position(step, 0));     // forward: (m, mbar) then (mbar, m); backward chains: the other way round
    L2HMC_TRY(position(step, 1));
    L2HMC_TRY(momentum(step));
  }

  // ---- accept probability, Metropolis-Hastings select [+ mod 2 pi] [+ observables]
  L2HMC_TRY(l2hmc_u1_hamiltonian(x, v, a->beta, N, T, X, h_new, st));
  L2HMC_TRY(l2hmc_accept_prob(h_old, h_new, sld, N, a->accept_prob, a->nonfinite_flag, st));
  if (a->x_out) {
    L2HMC_TRY(l2hmc_mh_select(a->x, x, nullptr, nullptr, nullptr, a->accept_prob, nullptr, nullptr, u_acc, N, D,
                              a->wrap_2pi, nullptr, nullptr, nullptr, a->x_out, a->accepted, st));
    if (a->obs_action || a->obs_plaq || a->obs_charge)
      L2HMC_TRY(l2hmc_u1_observables(a->x_out, N, T, X, a->obs_action, a->obs_plaq, nullptr, a->obs_charge, st));
  }
#undef L2HMC_TRY
  return L2HMC_OK;
}
