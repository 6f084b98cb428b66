/*
 * l2hmc_b200.h -- C ABI of the B200-native L2HMC / HMC leapfrog path for the 2D U(1)
 * lattice gauge model.
 *
 * The reference (saforem2/l2hmc) is pure Python on TensorFlow 1.x and has no FFI; the path
 * sits behind the Python classes GaugeLattice / GaugeDynamics / GenericNet / ConvNet3D.
 * The host-side mirror of those classes lives in l2hmc_b200/*.py and binds the entry
 * points below with ctypes (see INTEGRATION.md for the stub a maintainer would add to the
 * reference).  Each entry point names the reference code it replaces (file:line under
 * l2hmc/ in the reference tree).
 *
 * Conventions (all entry points):
 *   - every pointer is caller-owned DEVICE memory (fp32 unless stated), row-major;
 *     a chain is D = 2*T*X contiguous floats laid out [t][x][mu]
 *     (mu = 0 is the link along t, mu = 1 along x) == lattice.samples (N,T,X,2) flattened
 *     (lattice/lattice.py:121-129, 351-352);
 *   - nothing is allocated, nothing is synchronised: work is enqueued on `stream`
 *     (a cudaStream_t passed as void*; NULL = legacy default stream);
 *   - return value 0 = enqueued, non-zero = L2HMC_ERR_*; l2hmc_last_error() gives the text;
 *   - `nonfinite_flag` (nullable, device int32) is OR-ed with 1 when an exp() produced a
 *     non-finite value -- the device-side mirror of tf.check_numerics
 *     (dynamics/gauge_dynamics.py:26-28); the host wrapper raises on it;
 *   - re-entrant, no global state apart from the thread-local error string, the launch counter and the scratch
 *     array of l2hmc_col_sum_static.
 */
#ifndef L2HMC_B200_H_
#define L2HMC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define L2HMC_OK 0
#define L2HMC_ERR_INVALID 1     /* bad argument (null pointer, size, alignment) */
#define L2HMC_ERR_UNSUPPORTED 2 /* shape has no kernel instantiation */
#define L2HMC_ERR_CUDA 3        /* launch failed; see l2hmc_last_error() */

#define L2HMC_ABI_VERSION 2

int l2hmc_abi_version(void);
const char* l2hmc_last_error(void);
/* Number of kernels this library has launched in the process so far (every launch site adds to one relaxed atomic
 * counter): difference it around a region to count the region's launches.  No counterpart in the reference; it is what
 * bench.py reports as `gpu_launches`. */
unsigned long long l2hmc_launch_count(void);

/* Device limits the host side sizes its launches with. */
int l2hmc_device_info(int* sm_count, int* max_smem_per_block);

/* ------------------------------------------------------------------ K1: plaquette stencil
 * Wilson action S(x) = sum_{t,x} (1 - cos P) and/or force g = d(beta*S)/dx.
 *   P(t,x) = U0(t,x) - U1(t,x) - U0(t,x+1) + U1(t+1,x)   (periodic, left-associated)
 * replaces GaugeLattice.total_action (lattice/lattice.py:337-362) and
 * GaugeDynamics.grad_potential (dynamics/gauge_dynamics.py:698-709, tf.gradients of the above).
 * `action` and `force` are each nullable; any T, X >= 1. */
int l2hmc_u1_action_force(const float* x, float beta, int N, int T, int X,
                          float* action /*[N]*/, float* force /*[N,2TX]*/, void* stream);

/* ------------------------------------------------------------------ K4: observables
 * action, avg plaquette = sum cos P / (T*X), real-valued charge sum(proj P)/2pi, and the
 * integer charge floor(0.1 + sum(proj P)/2pi), proj(P) = P - 2pi floor((P+pi)/2pi).
 * replaces GaugeLattice.calc_plaq_observables (lattice/lattice.py:285-313) and
 * GaugeModel._calc_total_actions/_calc_avg_plaqs/_calc_top_charges (gauge_model.py:683-716).
 * All outputs nullable. */
int l2hmc_u1_observables(const float* x, int N, int T, int X, float* action, float* avg_plaq,
                         float* q_float, int32_t* q_int, void* stream);

/* H = beta*S(x) + 0.5*sum_d v^2 per chain in one pass over x and v.
 * GaugeDynamics.hamiltonian (dynamics/gauge_dynamics.py:691-696). */
int l2hmc_u1_hamiltonian(const float* x, const float* v, float beta, int N, int T, int X,
                         float* h /*[N]*/, void* stream);

/* 0.5 * sum_d v^2 per chain.  GaugeDynamics.kinetic_energy (dynamics/gauge_dynamics.py:683-689). */
int l2hmc_kinetic_energy(const float* v, int N, int D, float* ke, void* stream);

/* ------------------------------------------------------------------ K2: augmented leapfrog updates
 * dir = +1: GaugeDynamics._update_momentum_forward (dynamics/gauge_dynamics.py:486-508)
 *     v' = v*exp(eps/2*S) - eps/2*(exp(eps*Q)*g - T),      logdet += sum_d  eps/2*S
 * dir = -1: _update_momentum_backward (:537-561)
 *     v' = exp(-eps/2*S)*(v + eps/2*(exp(eps*Q)*g - T)),   logdet += sum_d -eps/2*S
 * S,T,Q nullable together (plain HMC, :102-108).  `logdet` [N] is ACCUMULATED into
 * (nullable).  v_out may alias v. */
int l2hmc_update_momentum(const float* v, const float* g, const float* S, const float* Tr,
                          const float* Q, float eps, int dir, int N, int D, float* v_out,
                          float* logdet, int32_t* nonfinite_flag, void* stream);

/* dir = +1: _update_position_forward (dynamics/gauge_dynamics.py:511-534)
 *     x' = m*x + mbar*(x*exp(eps*S) + eps*(exp(eps*Q)*v + T)),  logdet += sum_d mbar*eps*S
 * dir = -1: _update_position_backward (:565-590)
 *     x' = m*x + mbar*exp(-eps*S)*(x - eps*(exp(eps*Q)*v + T)), logdet += sum_d mbar*(-eps*S)
 * `mask` is one row [D] of GaugeDynamics.mask (:651-661); invert_mask != 0 swaps (m, mbar),
 * the second half-update of each LF step (:434-437).  No mod-2pi here (as in the reference). */
int l2hmc_update_position(const float* x, const float* v, const float* S, const float* Tr,
                          const float* Q, const float* mask, int invert_mask, float eps, int dir,
                          int N, int D, float* x_out, float* logdet, int32_t* nonfinite_flag,
                          void* stream);

/* out = mask_row (*) x  (or (1-mask_row) (*) x): the `mask * position` network input
 * (dynamics/gauge_dynamics.py:515-517). */
int l2hmc_mask_mul(const float* x, const float* mask, int invert_mask, int N, int D, float* out,
                   void* stream);

/* ------------------------------------------------------------------ accept / reject
 * GaugeDynamics._compute_accept_prob (dynamics/gauge_dynamics.py:592-609):
 *     p = exp(min(h_old - h_new + sumlogdet, 0)), non-finite -> 0. */
int l2hmc_accept_prob(const float* h_old, const float* h_new, const float* sumlogdet /*nullable*/,
                      int N, float* p, int32_t* nonfinite_flag, void* stream);

/* GaugeDynamics.apply_transition's blends (dynamics/gauge_dynamics.py:220-257):
 *   fm = [u_dir > 0.5]; x_post = fm*x_f + (1-fm)*x_b (same for v, p);
 *   am = [p > u_acc] (strict); x_out = am*x_post + (1-am)*x.
 * wrap_2pi != 0 additionally applies the host-side np.mod(x_out, 2pi) of
 * gauge_model.py:1180,1388 to x_out.  x_b/v_b/p_b nullable (single direction computed).
 * accepted [N] uint8 nullable. */
int l2hmc_mh_select(const float* x, const float* x_f, const float* x_b, const float* v_f,
                    const float* v_b, const float* p_f, const float* p_b, const float* u_dir,
                    const float* u_acc, int N, int D, int wrap_2pi, float* x_post, float* v_post,
                    float* p, float* x_out, uint8_t* accepted, void* stream);

/* ------------------------------------------------------------------ fused plain-HMC path (S=T=Q=0)
 * One leapfrog step per launch for all chains, x and v streamed once (16 B / link-update):
 *     v -= s*eps/2*g(x);  x += s*eps*v;  v -= s*eps/2*g(x)       s = dir[n] in {+1,-1}
 * == _forward_lf / _backward_lf with the hmc stubs (dynamics/gauge_dynamics.py:102-108,412-483).
 * `dir` [N] int8 nullable (NULL = all +1).  x_out/v_out may alias x/v.
 * Shapes: (T,X) in {4x4, 8x8, 16x16, 32x32, 8x16, 16x32}; others -> L2HMC_ERR_UNSUPPORTED
 * (the host falls back to K1 + K2). */
int l2hmc_u1_hmc_step(const float* x, const float* v, float beta, float eps, const int8_t* dir,
                      int N, int T, int X, float* x_out, float* v_out, void* stream);

/* Whole Monte-Carlo step in one launch: momentum (injected or Philox), M leapfrog steps with
 * links and momenta resident on-chip, both Hamiltonians, accept probability, direction and
 * accept/reject select == GaugeDynamics.apply_transition with hmc=True
 * (dynamics/gauge_dynamics.py:195-313), computing only the direction each chain selects.
 *
 *   momentum: v_f / v_b [N,D] are the draws of the forward / backward transition_kernel
 *             (:269); chain n uses v_f if it goes forward else v_b.  v_b NULL -> v_f for all.
 *             v_f NULL -> drawn in-kernel: Philox4x32-10 keyed (seed), counter
 *             (chain0 + n, link, mc_step) -- independent of sharding.
 *   direction: u_dir [N] nullable; NULL -> in-kernel uniform if v_f is NULL, else
 *             `default_dir` (+1/-1) for all chains (== transition_kernel(forward=...)).
 *   accept:   u_acc [N] nullable; NULL -> in-kernel uniform if v_f is NULL, else no select.
 *   outputs (each nullable): x_post, v_post [N,D], accept_prob [N], x_out [N,D],
 *             accepted [N] uint8, h_old / h_new [N]; obs_action / obs_plaq [N] and
 *             obs_charge [N] int32: the observables of l2hmc_u1_observables for the state
 *             each chain ends the step in (x_post if accepted, else x), evaluated on the
 *             un-wrapped links at no extra HBM traffic.
 */
typedef struct {
  const float* x;
  const float* v_f;
  const float* v_b;
  const float* u_dir;
  const float* u_acc;
  float beta;
  float eps;
  int num_steps;
  int default_dir;
  int wrap_2pi;
  uint64_t seed;
  uint64_t mc_step;
  uint64_t chain0;
  int N, T, X;
  float* x_post;
  float* v_post;
  float* accept_prob;
  float* x_out;
  uint8_t* accepted;
  float* h_old;
  float* h_new;
  int32_t* nonfinite_flag;
  float* obs_action;
  float* obs_plaq;
  int32_t* obs_charge;
  const uint64_t* mc_step_dev; /* nullable device counter ADDED to mc_step inside the kernel: a captured CUDA graph of
                                  the step replays with fresh draws when l2hmc_counter_add advances it in the graph */
} l2hmc_hmc_mc_args;

int l2hmc_u1_hmc_mc_step(const l2hmc_hmc_mc_args* args, void* stream);

/* The momenta / uniforms the in-kernel generator would use, written out (tests, and the
 * generic path's free-running mode).  normals [N,D], u_dir/u_acc [N], each nullable. */
int l2hmc_philox_draws(uint64_t seed, uint64_t mc_step, uint64_t chain0, int N, int D,
                       float* normals, float* u_dir, float* u_acc, void* stream);
/* Same with the MC step = mc_step + *mc_step_dev read on the device (mc_step_dev nullable), and the one-thread kernel
 * that advances such a counter in stream order: together they let a whole MC step -- every launch of
 * l2hmc_u1_l2hmc_mc_step / l2hmc_u1_hmc_mc_step is capturable, nothing allocates or synchronises -- be captured ONCE as
 * a CUDA graph and replayed per step (the small-batch configs are launch-latency-bound: host loop gauge_model.py:1158-1177,
 * 1379-1388). */
int l2hmc_philox_draws_dev(uint64_t seed, uint64_t mc_step, const uint64_t* mc_step_dev, uint64_t chain0, int N, int D,
                           float* normals, float* u_dir, float* u_acc, void* stream);
int l2hmc_counter_add(uint64_t* counter, uint64_t inc, void* stream);

/* ------------------------------------------------------------------ K3: S/T/Q networks
 * Dense trunk + heads of GenericNet.call (network/generic_net.py:129-146) and of ConvNet3D.call
 * after its conv front-end (network/conv_net.py:262-280):
 *     h1 = relu(in0.Wv + bv + in1.Wx + bx + tau.Wt + bt)     (the FIRST input goes through v_layer)
 *     h2 = relu(h1.Wh + bh)
 *     S  = tanh(h2.Ws + bs) * exp(coeff_s);  T = h2.Wt' + bt';  Q = (h2.Wq + bq) * exp(coeff_q)
 * Weights are device fp32 in Keras (in, out) layout, caller-owned.  `packed` (nullable) is the
 * buffer filled by l2hmc_net_pack(): weights pre-tiled and TF32-split for the tcgen05 path, used
 * when the shapes qualify (rows >= 128, K, H, D multiples of the tile sizes). */
typedef struct {
  int K;                          /* features of each of the two inputs */
  int H;                          /* hidden width */
  int D;                          /* outputs per head (= x_dim) */
  const float *w_in0, *b_in0;     /* v_layer: kernel [K,H], bias [H] (first input) */
  const float *w_in1, *b_in1;     /* x_layer: kernel [K,H], bias [H] (second input) */
  const float *w_t, *b_t;         /* t_layer: kernel [2,H], bias [H] */
  const float *w_h, *b_h;         /* h_layer: kernel [H,H], bias [H] */
  const float *w_s, *b_s;         /* scale_layer: kernel [H,D], bias [D] */
  const float *w_tr, *b_tr;       /* translation_layer */
  const float *w_q, *b_q;         /* transformation_layer */
  const float *coeff_s, *coeff_q; /* [D] */
  int tanh_on_q;                  /* 0 for the gauge nets; 1 = MoG net (utils/network.py:108-111) */
  const void* packed;             /* from l2hmc_net_pack(), nullable */
} l2hmc_net_weights;

#define L2HMC_NET_STQ 0      /* write S, T, Q */
#define L2HMC_NET_MOMENTUM 1 /* heads feed _update_momentum_{forward,backward} in the GEMM epilogue */
#define L2HMC_NET_POSITION 2 /* heads feed _update_position_{forward,backward} in the GEMM epilogue */

typedef struct {
  const float* in0;      /* [N,K] first trunk input */
  const float* in1;      /* [N,K] second trunk input */
  const float* in1_mask; /* nullable [K] 0/1 row multiplied into in1 on load (the `mask * position`
                            input of position_fn, dynamics/gauge_dynamics.py:515-517) */
  int in1_mask_invert;   /* use 1 - mask */
  float tau_cos, tau_sin;/* the time embedding row (dynamics/gauge_dynamics.py:625-633) */
  int N;
  int mode;              /* L2HMC_NET_* */
  /* L2HMC_NET_STQ: outputs [N,D] */
  float *S, *T, *Q;
  /* update modes (formulas of l2hmc_update_momentum / l2hmc_update_position) */
  float eps;
  int dir;               /* +1 / -1 */
  const float* x;        /* POSITION: links to update [N,D] */
  const float* v;        /* MOMENTUM: momenta to update; POSITION: momenta [N,D] */
  const float* g;        /* MOMENTUM: force [N,D] */
  const float* mask;     /* POSITION: mask row [D] */
  int mask_invert;       /* POSITION: swap (m, mbar) */
  float* out;            /* updated v (MOMENTUM) or x (POSITION), [N,D]; may alias the input */
  float* logdet;         /* [N], accumulated into; nullable */
  int32_t* nonfinite_flag;
  void* workspace;       /* >= l2hmc_net_workspace_bytes(...) bytes of device memory */
  size_t workspace_bytes;
  int path;              /* L2HMC_NET_PATH_*: AUTO picks tcgen05 when the shapes are dense contractions */
  /* Mixed-direction batches.  The reference integrates every chain forward AND backward and blends
     the two with exact 0/1 masks (dynamics/gauge_dynamics.py:211-241); here a chain runs only the
     direction it selected, and both kinds share one launch: row_set[n] in {0,1} picks, per chain,
     between the fields above (0) and the *_alt fields (1).  NULL: every chain uses set 0. */
  const uint8_t* row_set;      /* nullable [N] */
  float tau_cos_alt, tau_sin_alt;
  const float* in1_mask_alt;   /* nullable [K] */
  int in1_mask_invert_alt;
  int dir_alt;                 /* +1 / -1 */
  const float* mask_alt;       /* POSITION: mask row [D] */
  int mask_invert_alt;
  /* Training tape: when given, the hidden activations h1 / h2 [N,H] (after the ReLU, row-major fp32) are written
     here and survive the call, for the backward pass (l2hmc_gemm / l2hmc_update_bwd).  Nullable, together. */
  float *h1_out, *h2_out;
} l2hmc_net_call;

#define L2HMC_NET_PATH_AUTO 0
#define L2HMC_NET_PATH_FFMA 1
#define L2HMC_NET_PATH_TCGEN05 2 /* L2HMC_ERR_UNSUPPORTED if the shape / missing `packed` rules it out */

size_t l2hmc_net_workspace_bytes(int K, int H, int D, int N);
int l2hmc_net_run(const l2hmc_net_weights* w, const l2hmc_net_call* call, void* stream);

/* Pre-tiled, TF32-split copy of the dense weights for the tcgen05 path (3xTF32: hi.lo + lo.hi + hi.hi,
 * fp32 accumulation in TMEM).  `packed` is caller-owned device memory of l2hmc_net_pack_bytes(K,H,D)
 * bytes; re-pack after the weights change.  w->packed is ignored by l2hmc_net_pack itself. */
size_t l2hmc_net_pack_bytes(int K, int H, int D);
/* Smallest batch N for which L2HMC_NET_PATH_AUTO takes the tcgen05 layers for a (K, H, D) network (INT_MAX when the
 * shape never does).  Below it the chunked FFMA layers run, whose result for a chain does not depend on the batch it
 * was launched with; a caller that splits a batch (l2hmc_b200/sampler.py) keeps every part on the same side of this
 * number so that splitting does not change a bit.  No reference counterpart (dispatch detail of this library). */
int l2hmc_net_tc_min_rows(int K, int H, int D);
int l2hmc_net_pack(const l2hmc_net_weights* w, void* packed, size_t packed_bytes, void* stream);

/* Diagnostic: C[M,Nc] = relu(A[M,K] . W[K,Nc]) through the tcgen05 layer kernel and its operand
 * converters (row-major fp32 in / out; bn = 128 or 256 selects the N tile). */
size_t l2hmc_tc_selftest_bytes(int M, int Nc, int K);
int l2hmc_tc_selftest(const float* A, const float* W, int M, int Nc, int K, int bn, float* C, void* scratch,
                      void* stream);

/* Conv front-end of ConvNet3D for one input stream (network/conv_net.py:247-260, reshape_5D
 * :283-311; channels_last): (N,T,X,2,1) -> Conv3D(F,(3,3,2),same,relu) -> MaxPool3D(2,2,same)
 * -> Conv3D(2F,(2,2,2),same,relu) -> MaxPool3D(2,2,same) -> flatten, written as
 * feat [N, l2hmc_conv3d_num_features(T,X,F)] (the trunk's in0 / in1).  Kernels in Keras layout
 * (kd,kh,kw,cin,cout).  `mask` (nullable [2TX]) multiplies the input on load (position_fn's
 * `mask * x`). */
typedef struct {
  const float *w1, *b1; /* conv_{x,v}1: kernel [3,3,2,1,F], bias [F] */
  const float *w2, *b2; /* conv_{x,v}2: kernel [2,2,2,F,2F], bias [2F] */
  int F;
} l2hmc_conv3d_weights;

int l2hmc_conv3d_num_features(int T, int X, int F);
int l2hmc_conv3d_features(const l2hmc_conv3d_weights* w, const float* in, const float* mask, int mask_invert,
                          int N, int T, int X, float* feat, void* stream);
/* Same, for mixed-direction batches: chain n uses (mask_alt, mask_invert_alt) when row_set[n] != 0
 * (see l2hmc_net_call.row_set); row_set == NULL is l2hmc_conv3d_features. */
int l2hmc_conv3d_features_sets(const l2hmc_conv3d_weights* w, const float* in, const float* mask, int mask_invert,
                               const float* mask_alt, int mask_invert_alt, const uint8_t* row_set, int N, int T,
                               int X, float* feat, void* stream);

/* Routing tape of the conv front-end for the training pass: what the adjoint needs of the forward pass, 5.5 KB per chain
 * and call at 16x16 instead of recomputing both convolutions (tf.gradients keeps the activations themselves).
 *   p1   [N][T1*X1*F]   pooled conv-1 output (after bias + ReLU), T1 = ceil(T/2), X1 = ceil(X/2)
 *   arg1 [N][T1*X1*F]   uint8: which of the window's (dt, dx, mu) candidates was the maximum, dt*4 + dx*2 + mu (the
 *                       first one in that order), or 255 when the pooled value is 0 (ReLU-dead: no gradient)
 *   arg2 [N][features]  uint8: dt*2 + dx of the second pool's maximum, or 255 when dead
 * All three caller-owned device memory, written by l2hmc_conv3d_features_tape, read by l2hmc_conv3d_backward. */
typedef struct {
  float* p1;
  uint8_t* arg1;
  uint8_t* arg2;
} l2hmc_conv3d_tape;

int l2hmc_conv3d_tape_elems(int T, int X, int F); /* T1*X1*F: elements of p1 / arg1 per chain */
/* l2hmc_conv3d_features_sets that also records the tape (tape non-null). */
int l2hmc_conv3d_features_tape(const l2hmc_conv3d_weights* w, const float* in, const float* mask, int mask_invert,
                               const float* mask_alt, int mask_invert_alt, const uint8_t* row_set, int N, int T, int X,
                               float* feat, const l2hmc_conv3d_tape* tape, void* stream);

/* ------------------------------------------------------------------ one host call = one L2HMC MC step
 * GaugeDynamics.apply_transition WITH the S/T/Q networks (dynamics/gauge_dynamics.py:195-313, 412-483): momentum and
 * direction draws, M augmented leapfrog steps (each chain in the direction it drew, both kinds in the same
 * launches), accept probability, Metropolis-Hastings select [+ np.mod 2 pi, gauge_model.py:1180,1388]
 * [+ observables of the new state, gauge_model.py:683-716].  Only launches kernels of this library, in stream
 * order; nothing is allocated or synchronised.  v_f == NULL: momenta and uniforms come from the Philox stream
 * (seed, chain0 + n, mc_step), the same numbers l2hmc_philox_draws returns.  Generic nets: the four conv pointers
 * NULL.  Conv nets: xconv_in0 / xconv_in1 are position_fn's conv stacks for its first / second input (conv_v*,
 * conv_x*), vconv_* momentum_fn's; the nets' K is then the conv feature count. */
typedef struct {
  int N, T, X, M;
  float eps, beta;
  const float* masks;   /* [M][2TX] device */
  const float* ts;      /* [M][2] HOST array: (cos, sin)(2 pi t / M) rows, dynamics/gauge_dynamics.py:625-633 */
  const l2hmc_net_weights* xnet; /* position_fn */
  const l2hmc_net_weights* vnet; /* momentum_fn */
  const l2hmc_conv3d_weights *xconv_in0, *xconv_in1, *vconv_in0, *vconv_in1; /* nullable (all or none) */
  int net_path;         /* L2HMC_NET_PATH_* */
  const float* x;       /* [N,2TX] */
  const float *v_f, *v_b;     /* [N,2TX] injected momenta, nullable (v_b NULL: v_f both ways) */
  const float *u_dir, *u_acc; /* [N] injected uniforms (required with v_f) */
  uint64_t seed, mc_step, chain0;
  int wrap_2pi;
  float *x_post, *v_post; /* [N,2TX] proposal (also the in-place working state) */
  float* accept_prob;     /* [N] */
  float* x_out;           /* [N,2TX] nullable */
  uint8_t* accepted;      /* [N] nullable */
  float *obs_action, *obs_plaq; /* [N] nullable: observables of x_out */
  int32_t* obs_charge;
  int32_t* nonfinite_flag;
  void* workspace;        /* >= l2hmc_u1_l2hmc_mc_workspace_bytes(args), 256-byte aligned */
  size_t workspace_bytes;
  const uint64_t* mc_step_dev; /* nullable device counter added to mc_step on the device (see l2hmc_hmc_mc_args) */
} l2hmc_l2hmc_mc_args;

size_t l2hmc_u1_l2hmc_mc_workspace_bytes(const l2hmc_l2hmc_mc_args* a);
int l2hmc_u1_l2hmc_mc_step(const l2hmc_l2hmc_mc_args* a, void* stream);

/* ------------------------------------------------------------------ MoG / generic-target path (config 5)
 * Gaussian-mixture target of utils/distributions.py:124-158:
 *     E(x) = -logsumexp_i( -0.5 (x-mu_i)^T S_i (x-mu_i) + log c_i ) / temperature
 * computed row-wise (the reference forms an N x N product and takes its diagonal, :32-39).
 * mus [m,k], i_sigmas [m,k,k] (= inv(sigma_i)), log_consts [m] (= log(pi_i / sqrt((2pi)^k det sigma_i))),
 * all device fp32.  k <= 8, m <= 16. */
typedef struct {
  int k, m;
  const float* mus;
  const float* i_sigmas;
  const float* log_consts;
} l2hmc_gmm;

/* energy [N] and force = dE/dx [N,k] (each nullable): Dynamics.energy / grad_energy
 * (utils/dynamics.py:227-242). */
int l2hmc_gmm_energy_force(const l2hmc_gmm* gmm, const float* x, float temperature, int N, float* energy,
                           float* force, void* stream);

/* One MC step of the MoG sampler in ONE launch, one thread per chain, both networks' weights in
 * shared memory: utils/sampler.py:28-59 propose(do_mh_step=True) over
 * Dynamics.forward / backward (utils/dynamics.py:120-319), integrating only the direction each
 * chain selects.  Nets are utils/network.py:89-114 (tanh on both S and F heads): pass
 * l2hmc_net_weights with K = D = k, H <= 64, tanh_on_q = 1 (xnet / vnet NULL when hmc != 0).
 *   v_f / v_b [N,k]: momentum draws of forward / backward (v_b NULL -> v_f); v_f NULL -> Philox
 *   dir_mask [N]: 1 = forward, 0 = backward (the reference's random {0,1} mask); NULL -> Philox
 *                 uniform >= 0.5, or all forward when momenta are injected; ignored (forward) for hmc
 *   u_acc [N]: accept if p - u >= 0; NULL -> Philox when v_f is NULL, else no accept step
 *   mask [L,k]: Dynamics.mask (:85-100)
 * outputs (each nullable): Lx, Lv [N,k], px [N], x_out [N,k], accepted [N] uint8, mode [N] int32 =
 * index of the mixture mean nearest to x_out (utils/trajectories.py:39-60, for the tunneling rate). */
typedef struct {
  const float* x;
  const float* v_f;
  const float* v_b;
  const float* dir_mask;
  const float* u_acc;
  const float* mask;
  float eps;
  float temperature;
  int L;
  int hmc;
  int N;
  uint64_t seed, mc_step, chain0;
  const l2hmc_net_weights* xnet;
  const l2hmc_net_weights* vnet;
  float* Lx;
  float* Lv;
  float* px;
  float* x_out;
  uint8_t* accepted;
  int32_t* mode;
  int32_t* nonfinite_flag;
} l2hmc_mog_args;

int l2hmc_mog_mc_step(const l2hmc_gmm* gmm, const l2hmc_mog_args* args, void* stream);

/* Value of the MoG training loss, mog_model.py:324-355 (GaussianMixtureModel._create_loss):
 *   v1 = |x - Lx|^2 * px + 1e-4,  v2 = |z - Lz|^2 * pz + 1e-4      (squared jump distances, per chain)
 *   loss = scale * (mean 1/v1 + mean 1/v2) + (-mean v1 - mean v2) / scale
 * with (Lx, px) = propose(x) and (Lz, pz) = propose(z ~ N(0,1)) from l2hmc_mog_mc_step.  x, Lx, z, Lz [N,k]; px, pz
 * [N]; v1, v2 [N] nullable outputs; loss [1].  Two launches, fixed summation order (deterministic).  The VALUE only:
 * the reference's MoG driver cannot run as shipped (mog_model.py:51) and SURVEY 8(f) does not ask for its gradient. */
size_t l2hmc_mog_loss_workspace_bytes(void);
int l2hmc_mog_loss(const float* x, const float* Lx, const float* px, const float* z, const float* Lz, const float* pz,
                   long long N, int k, float scale, float* v1, float* v2, float* loss, void* workspace,
                   size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ loss value, run statistics
 * Training loss VALUE of gauge_model.py:728-803 (_calc_loss) given the two chains' transitions
 * (x --dynamics--> x_prop, px, x_out;  z ~ N(0,1) --dynamics--> pz): per chain
 *   x_std = sum_d metric(x, x_prop) * px + 1e-3,  z_std = aux * (sum_d metric(z, x_prop) * pz + 1e-3)
 *   std   = (ls * (1/x_std + 1/z_std) - (x_std + z_std) / ls) * std_weight
 *   chg   = charge_weight * ((px |Qfft(x) - Qfft(x_prop)| + 1e-3) + aux * (pz |Qfft(z) - Qfft(x_prop)| + 1e-3))
 * with Qfft = sum project_angle_approx(P) / 2pi (gauge_model.py:94-108, harmonics 1..4); loss[0] = mean over the
 * chains (fixed summation tree); x_dq[n] = int32(|Q(x) - Q(x_out)|) with the exact projection (:761-762).
 * The backward pass (tf.gradients, :805-832) is not part of this library yet. */
#define L2HMC_METRIC_L1 0
#define L2HMC_METRIC_L2 1
#define L2HMC_METRIC_COS 2
#define L2HMC_METRIC_COS2 3
#define L2HMC_METRIC_COS_DIFF 4 /* gauge_model.py:631-657 */

typedef struct {
  int metric; /* L2HMC_METRIC_* */
  float loss_scale, aux_weight, std_weight, charge_weight;
} l2hmc_loss_params;

int l2hmc_u1_loss(const l2hmc_loss_params* prm, const float* x, const float* x_prop, const float* z,
                  const float* x_out, const float* px, const float* pz, int N, int T, int X,
                  float* per_chain /* [N] */, int32_t* x_dq /* [N], nullable */, float* loss /* [1], nullable */,
                  void* stream);
/* _calc_top_charges(x, fft=True), gauge_model.py:699-713: q [N] float. */
int l2hmc_u1_top_charges_fft(const float* x, int N, int T, int X, float* q, void* stream);

/* Streaming form of calc_observables_stats (gauge_model.py:1473-1533).  sum / sumsq: [4][N] fp64 zero-initialised by
 * the caller, rows = action, plaquette, charge, charge^2; hist: [2*qmax + 2] zero-initialised counts of the charges
 * -qmax..qmax (+ one overflow slot), nullable.  Per MC step call l2hmc_stats_update with counted = 0 for the
 * thermalisation steps (the first num_steps // therm_frac; only charge^2 accumulates there, the reference squares
 * before it cuts) and 1 afterwards; charges are truncated toward zero like np.array(..., dtype=int).
 * l2hmc_stats_finalize writes mean and scipy.stats.sem (ddof = 1), both [4][N] fp64. */
typedef struct {
  int N;
  int qmax;
  double* sum;
  double* sumsq;
  unsigned long long* hist;
} l2hmc_run_stats;

int l2hmc_stats_update(const l2hmc_run_stats* s, const float* action, const float* plaq, const float* charge,
                       int counted, void* stream);
int l2hmc_stats_finalize(const l2hmc_run_stats* s, long long n_counted, long long n_total, double* mean, double* err,
                         void* stream);
/* utils/trajectories.py:63-95: switches[n] += (mode[n] != prev[n]) when have_prev; prev[n] = mode[n]. */
int l2hmc_mode_switches(const int32_t* mode, int32_t* prev, uint32_t* switches, int have_prev, int N, void* stream);

/* ------------------------------------------------------------------ training: backward pass + optimiser
 * SURVEY section 8(f) ranks 1-2.  The reference obtains every gradient from tf.gradients(loss, dynamics.variables)
 * (gauge_model.py:805-832); here each adjoint is an explicit kernel and the host (l2hmc_b200/training.py) walks the
 * tape the forward pass recorded.  Everything is deterministic (no atomics; split-K sums in fixed order). */

/* fp32 GEMM with general strides (NN / NT / TN without copies), up to three K segments summed into one product:
 *     C[m,n] (=|+=) scale[n] * gate[m,n] * sum_seg sum_k A_seg(m,k) * B_seg(k,n)
 *   A_seg(m,k) = A[m*a_rs + k*a_cs],  B_seg(k,n) = B[k*b_rs + n*b_cs]
 *   relu_gate (nullable): the product is zeroed where gate[m*ldg + n] <= 0   (d relu: dZ = dH * [h > 0])
 *   col_scale (nullable [N]): multiplies column n by scale[n] (or 1 - scale[n]); rows with row_set[m] != 0 use
 *   col_scale_alt / col_scale_invert_alt  (the `mask * x` input of position_fn, dynamics/gauge_dynamics.py:515-517)
 * ksplit > 1 splits K (of every segment) over grid.z; the slices are summed in slice order and the epilogue applied
 * once: up to 8 slices inside a thread-block cluster (rank 0 reads the other ranks' accumulators over distributed shared
 * memory), more through `workspace` (l2hmc_sgemm_workspace_bytes) and a second kernel -- the same sums either way.
 * Used by the deferred weight-gradient GEMMs, whose K is the whole tape, and -- ksplit = 0, chosen by the library --
 * the few-tile per-call GEMMs of a small batch.  M <= 4 over K >= 2048 with row-major B (the t_layer kernel's
 * gradient) runs as a weighted column sum with the library's own slab count. */
typedef struct {
  const float* A;
  long long a_rs, a_cs;
  const float* B;
  long long b_rs, b_cs;
  int K;
} l2hmc_gemm_seg;

typedef struct {
  int M, N, nseg;
  l2hmc_gemm_seg seg[3];
  float* C;
  long long ldc;
  int accumulate;
  const float* relu_gate;
  long long ldg;
  const float* col_scale;
  int col_scale_invert;
  const uint8_t* row_set;
  const float* col_scale_alt;
  int col_scale_invert_alt;
} l2hmc_gemm_args;

size_t l2hmc_sgemm_workspace_bytes(const l2hmc_gemm_args* g, int ksplit);
int l2hmc_sgemm(const l2hmc_gemm_args* g, int ksplit, void* workspace, size_t workspace_bytes, void* stream);
/* The same product on the tcgen05 tensor cores (3xTF32 split: hi.lo + lo.hi + hi.hi, fp32 accumulation flushed every
 * 128 K into round-to-nearest register sums, so the result is fp32-accurate): both operands are first converted from
 * their strided fp32 form into the tiled TS32 layout inside `workspace` (l2hmc_tc_gemm_workspace_bytes), any of the
 * NN / NT / TN forms without a transposed copy; same epilogues as l2hmc_sgemm, written row-major.  Long-K products
 * with few output tiles are split over K (fixed ranges, fixed summation order) so that all SMs take part. */
size_t l2hmc_tc_gemm_workspace_bytes(const l2hmc_gemm_args* g);
int l2hmc_tc_gemm(const l2hmc_gemm_args* g, void* workspace, size_t workspace_bytes, void* stream);
/* What the training step calls: l2hmc_tc_gemm when the product is a dense contraction (M >= 128, N >= 64, K >= 128 and
 * M*N*K >= 2.5e8), l2hmc_sgemm(ksplit) otherwise. */
size_t l2hmc_gemm_workspace_bytes(const l2hmc_gemm_args* g, int ksplit);
int l2hmc_gemm(const l2hmc_gemm_args* g, int ksplit, void* workspace, size_t workspace_bytes, void* stream);

/* The four sub-updates (formulas of l2hmc_update_momentum / l2hmc_update_position) with a per-chain parameter set:
 * chains with row_set[n] != 0 use dir_alt / mask_alt / invert_alt (mixed-direction batches).  S, T, Q nullable
 * together (plain HMC).  `logdet` [N] is accumulated into. */
int l2hmc_train_update_fwd(int position, const float* x, const float* v, const float* g, const float* S,
                           const float* Tr, const float* Q, const float* mask, int invert, const float* mask_alt,
                           int invert_alt, const uint8_t* row_set, float eps, int dir, int dir_alt, int N, int D,
                           float* out, float* logdet, int32_t* nonfinite_flag, void* stream);
/* out[n,:] = m (*) x[n,:] with the chain's mask set (materialised network input for the tape). */
int l2hmc_mask_mul_sets(const float* x, const float* mask, int invert, const float* mask_alt, int invert_alt,
                        const uint8_t* row_set, int N, int D, float* out, void* stream);

/* Adjoint of one sub-update (dynamics/gauge_dynamics.py:486-590).  In: the update's inputs from the tape, the
 * upstream gradients gv (momentum: d loss / d v') or gx (position: d loss / d x') and gld [N] = d loss / d sumlogdet.
 * Out (in place): momentum: gv <- d loss / d v;  dg [N,D] = the direct part of d loss / d force.
 *                 position: gx <- direct part of d loss / d x;  gv += d loss / d v (direct part).
 * With networks: dA [N,3D] = gradients of the three heads' pre-activations (scale | translation | transformation),
 * ps / pq [N,D] = dS*S / dQ*Q (their column sums are d loss / d coeff_scale / coeff_transformation);
 * deps [N] += per-chain d loss / d eps. */
typedef struct {
  int position; /* 0 = momentum update, 1 = position update */
  int N, D;
  float eps;
  int dir, dir_alt;
  const uint8_t* row_set;
  const float *x, *v, *g;
  const float *S, *T, *Q;
  const float *coeff_s, *coeff_q;
  const float *mask, *mask_alt;
  int invert, invert_alt;
  float *gx, *gv;
  const float* gld;
  float *dA, *dg, *ps, *pq, *deps;
  const float* eps_dev;   /* nullable: the step size read from device memory instead of `eps` (see l2hmc_train_args) */
} l2hmc_update_bwd_args;
int l2hmc_update_bwd(const l2hmc_update_bwd_args* a, void* stream);

/* out += d/dx [ force(x) . w ]  (force = d(beta*S)/dx): the Hessian-vector product that tf.gradients builds when it
 * differentiates through grad_potential (dynamics/gauge_dynamics.py:698-709), i.e. the network input `grad`. */
int l2hmc_u1_force_hvp(const float* x, const float* w, float beta, int N, int T, int X, float* out, void* stream);

/* a[n] = d loss / d (h_old - h_new + sumlogdet) given g_p = d loss / d accept_prob (dynamics/gauge_dynamics.py:592-609). */
int l2hmc_accept_prob_bwd(const float* h_old, const float* h_new, const float* sumlogdet, const float* g_p, int N,
                          float* a, void* stream);
/* Start of a trajectory's backward sweep: gx = gx_post - a * force(x_M), gv = -a * v_M  (gx_post nullable = 0). */
int l2hmc_traj_bwd_init(const float* a, const float* force, const float* v, const float* gx_post, int N, int D,
                        float* gx, float* gv, void* stream);

/* d loss / d (x_prop, px, pz) of l2hmc_u1_loss's mean loss (gauge_model.py:762-801). */
int l2hmc_u1_loss_bwd(const l2hmc_loss_params* prm, const float* x, const float* x_prop, const float* z,
                      const float* px, const float* pz, int N, int T, int X, float* g_xprop /* [N,2TX] */,
                      float* g_px /* [N] */, float* g_pz /* [N] */, void* stream);

/* Backward pass of the conv front-end of ConvNet3D for one input stream (adjoint of l2hmc_conv3d_features[_sets];
 * network/conv_net.py:247-260 under tf.gradients, gauge_model.py:823).  The forward activations are recomputed
 * on-chip from `in`.  g_feat [N, l2hmc_conv3d_num_features] is d loss / d features; g_in [N,2TX] is ACCUMULATED into
 * (+= mask (*) d loss / d masked input); `partial` [l2hmc_conv3d_bwd_rows(N)][l2hmc_conv3d_bwd_cols(F)] is
 * ACCUMULATED into: each row holds (d w1 [3,3,2,1,F] | d b1 [F] | d w2 [2,2,2,F,2F] | d b2 [2F]) of the chains
 * n = row mod rows; the caller zeroes it before the first call of a step and sums its rows (fixed order) at the end.
 * L2HMC_ERR_UNSUPPORTED when a chain's activations do not fit shared memory (32x32 with F = 32 WITHOUT a tape; with
 * the tape that shape fits). */
typedef struct {
  const l2hmc_conv3d_weights* w;
  const float* in;          /* [N,2TX] the stream's input (before masking) */
  const float* mask;        /* nullable [2TX] */
  int mask_invert;
  const float* mask_alt;    /* rows with row_set != 0 */
  int mask_invert_alt;
  const uint8_t* row_set;   /* nullable [N] */
  int N, T, X;
  const float* g_feat;
  float* g_in;
  float* partial;
  const l2hmc_conv3d_tape* tape; /* nullable: the forward pass's routing tape (l2hmc_conv3d_features_tape).  With it
                                    nothing is recomputed and every product runs over the NON-ZEROS of the pooled
                                    gradients (one per pooling window); without it both convolutions are recomputed
                                    on-chip.  Same gradients up to summation order. */
} l2hmc_conv3d_bwd_args;
int l2hmc_conv3d_bwd_rows(int N);
int l2hmc_conv3d_bwd_cols(int F);
int l2hmc_conv3d_backward(const l2hmc_conv3d_bwd_args* a, void* stream);

/* out[c] = sum_r B[r*ld + c] (bias / coeff_* gradients: batch sums of an activation gradient), two stages with a fixed
 * order; workspace: l2hmc_col_sum_workspace_bytes(rows, cols). */
size_t l2hmc_col_sum_workspace_bytes(int rows, int cols);
int l2hmc_col_sum(const float* B, long long ld, int rows, int cols, float* out, void* workspace, size_t workspace_bytes,
                  void* stream);

/* l2hmc_col_sum WITHOUT a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating library
 * call, here or on first use): the first-stage partials live in a file-scope `static __device__` array inside the
 * library (4 MiB; every GPU that loads the library holds it until the process exits).  No counterpart in the
 * reference: added for callers that cannot allocate (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernels, same summation order and therefore
 * the same bits as l2hmc_col_sum.  Limit: colsum row slabs x cols = min(512, max(1, rows / 32)) * cols <=
 * L2HMC_COL_SUM_STATIC_MAX_FLOATS, else L2HMC_ERR_INVALID with the sizes in l2hmc_last_error().  The device address
 * of the array is looked up once per device (cudaGetSymbolAddress, with that device current) on the first call
 * there -- make that first call outside stream capture.  All calls share the one array: calls that overlap on the
 * GPU (different streams, no ordering between them) race; calls on one stream, or ordered by events, do not. */
#define L2HMC_COL_SUM_STATIC_MAX_FLOATS (1 << 20)
int l2hmc_col_sum_static(const float* B, long long ld, int rows, int cols, float* out, void* stream);

/* out[0] = sum v[i] (square = 0) or sum v[i]^2 (square = 1), fixed two-stage tree; scratch: 1024 floats. */
int l2hmc_reduce_sum(const float* v, long long n, int square, float* out, float* scratch, void* stream);

/* tf.train.AdamOptimizer.apply_gradients over a flat parameter arena (gauge_model.py:934-945): lr_t = lr *
 * sqrt(1 - beta2^t) / (1 - beta1^t) is folded by the caller.  grad_scale multiplies the gradient (1 / world size after
 * a SUM all-reduce); grad_sumsq (nullable, device [1]) + clip_norm > 0 apply tf.clip_by_global_norm (:826-827). */
int l2hmc_adam_step(float* param, const float* grad, float* m, float* v, long long n, float lr_t, float beta1,
                    float beta2, float epsilon, float grad_scale, const float* grad_sumsq, float clip_norm,
                    void* stream);

/* tf.clip_by_global_norm (gauge_model.py:826-827) on the flat gradient arena, in place: grad *= clip_norm /
 * max(sqrt(grad_sumsq[0]), clip_norm), grad_sumsq from l2hmc_reduce_sum(square = 1).  The reference clips inside
 * _calc_loss_and_grads, i.e. per rank BEFORE Horovod averages (gauge_model.py:941-943); same bits as handing the clip
 * to l2hmc_adam_step. */
int l2hmc_clip_by_global_norm(float* grad, long long n, const float* grad_sumsq, float clip_norm, void* stream);

/* ------------------------------------------------------------------ one host call = loss + gradients of one step
 * GaugeModel._calc_loss_and_grads (gauge_model.py:805-832 over :728-803 and dynamics/gauge_dynamics.py:195-313):
 * both chains' transitions recorded on a tape, the loss, and the explicit backward sweep -- the launches of the
 * entry points above, in a fixed order, on a caller-provided workspace; nothing is allocated or synchronised.
 * Rows: the x-chain (rows 0..N-1) and the auxiliary z-chain (rows N..2N-1) are stacked; v_f / v_b [2N,D] and
 * u_dir [2N] are the draws of both (the reference's tf.random_normal / random_uniform, :754-758 over
 * dynamics/gauge_dynamics.py:211-248), u_acc [N] the x-chain's accept draw.  xnet == NULL: plain HMC (only g_eps).
 * Gradients are WRITTEN (not accumulated) through the l2hmc_net_grads / l2hmc_conv3d_grads pointer sets, which
 * mirror the weight structs.  forward_only != 0 stops after the loss. */
typedef struct {
  float *w_in0, *b_in0, *w_in1, *b_in1, *w_t, *b_t, *w_h, *b_h, *w_s, *b_s, *w_tr, *b_tr, *w_q, *b_q, *coeff_s, *coeff_q;
} l2hmc_net_grads;
typedef struct {
  float *w1, *b1, *w2, *b2;
} l2hmc_conv3d_grads;

typedef struct {
  int N, T, X, M;
  float eps, beta;
  const float* masks; /* [M][2TX] device */
  const float* ts;    /* [M][2] HOST (cos, sin) rows */
  const l2hmc_net_weights *xnet, *vnet;                                         /* nullable together */
  const l2hmc_conv3d_weights *xconv_in0, *xconv_in1, *vconv_in0, *vconv_in1;    /* nullable together */
  const float *x, *z;     /* [N,2TX] */
  const float *v_f, *v_b; /* [2N,2TX] */
  const float* u_dir;     /* [2N] */
  const float* u_acc;     /* [N] */
  l2hmc_loss_params loss;
  int forward_only;
  float* loss_out;        /* [1] */
  float* per_chain;       /* [N] */
  float* x_out;           /* [N,2TX] */
  float* px;              /* [N] */
  float* pz;              /* [N] nullable */
  float* x_prop;          /* [N,2TX] nullable */
  int32_t* x_dq;          /* [N] nullable */
  uint8_t* accepted;      /* [N] nullable */
  float* g_eps;           /* [1] */
  const l2hmc_net_grads *xnet_g, *vnet_g;
  const l2hmc_conv3d_grads *xconv_in0_g, *xconv_in1_g, *vconv_in0_g, *vconv_in1_g;
  int32_t* nonfinite_flag;
  void* workspace;        /* >= l2hmc_u1_train_workspace_bytes(args), 256-byte aligned */
  size_t workspace_bytes;
  void* first_bucket_event; /* nullable cudaEvent_t, recorded on `stream` as soon as g_eps and position_fn's gradients
                               (xnet_g, xconv_*_g) are complete, BEFORE momentum_fn's weight-gradient products are
                               enqueued: a data-parallel caller starts averaging that bucket on a side stream under the
                               rest of the backward pass (hvd.DistributedOptimizer overlaps the same way,
                               gauge_model.py:941-943) */
  const float* eps_dev;   /* nullable [1]: the trainable step size where the optimiser keeps it.  When given, every
                             kernel reads it from there and `eps` is ignored, so the caller need not copy it to the
                             host (and synchronise) after each optimiser step as the reference's session.run fetch of
                             dynamics.eps does (gauge_model.py:1166) */
} l2hmc_train_args;

size_t l2hmc_u1_train_workspace_bytes(const l2hmc_train_args* a);
int l2hmc_u1_train_grads(const l2hmc_train_args* a, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* L2HMC_B200_H_ */
