"""ctypes binding of the C-ABI CUDA library (include/l2hmc_b200.h).

There is deliberately NO fallback: if libl2hmc_b200.so cannot be loaded, or no CUDA device is
present, every compute entry point raises.  PyTorch is used only for device memory, streams and
torch.distributed.
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_longlong, c_size_t, c_uint64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libl2hmc_b200.so")

_lib = None


class HmcMcArgs(Structure):
    """l2hmc_hmc_mc_args (include/l2hmc_b200.h)."""
    _fields_ = [
        ("x", c_void_p), ("v_f", c_void_p), ("v_b", c_void_p), ("u_dir", c_void_p), ("u_acc", c_void_p),
        ("beta", c_float), ("eps", c_float), ("num_steps", c_int), ("default_dir", c_int), ("wrap_2pi", c_int),
        ("seed", c_uint64), ("mc_step", c_uint64), ("chain0", c_uint64),
        ("N", c_int), ("T", c_int), ("X", c_int),
        ("x_post", c_void_p), ("v_post", c_void_p), ("accept_prob", c_void_p), ("x_out", c_void_p),
        ("accepted", c_void_p), ("h_old", c_void_p), ("h_new", c_void_p), ("nonfinite_flag", c_void_p),
        ("obs_action", c_void_p), ("obs_plaq", c_void_p), ("obs_charge", c_void_p), ("mc_step_dev", c_void_p),
    ]


class NetWeights(Structure):
    """l2hmc_net_weights (include/l2hmc_b200.h)."""
    _fields_ = [
        ("K", c_int), ("H", c_int), ("D", c_int),
        ("w_in0", c_void_p), ("b_in0", c_void_p), ("w_in1", c_void_p), ("b_in1", c_void_p),
        ("w_t", c_void_p), ("b_t", c_void_p), ("w_h", c_void_p), ("b_h", c_void_p),
        ("w_s", c_void_p), ("b_s", c_void_p), ("w_tr", c_void_p), ("b_tr", c_void_p),
        ("w_q", c_void_p), ("b_q", c_void_p), ("coeff_s", c_void_p), ("coeff_q", c_void_p),
        ("tanh_on_q", c_int), ("packed", c_void_p),
    ]


class NetCall(Structure):
    """l2hmc_net_call (include/l2hmc_b200.h)."""
    _fields_ = [
        ("in0", c_void_p), ("in1", c_void_p), ("in1_mask", c_void_p), ("in1_mask_invert", c_int),
        ("tau_cos", c_float), ("tau_sin", c_float), ("N", c_int), ("mode", c_int),
        ("S", c_void_p), ("T", c_void_p), ("Q", c_void_p),
        ("eps", c_float), ("dir", c_int),
        ("x", c_void_p), ("v", c_void_p), ("g", c_void_p), ("mask", c_void_p), ("mask_invert", c_int),
        ("out", c_void_p), ("logdet", c_void_p), ("nonfinite_flag", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t), ("path", c_int),
        ("row_set", c_void_p), ("tau_cos_alt", c_float), ("tau_sin_alt", c_float),
        ("in1_mask_alt", c_void_p), ("in1_mask_invert_alt", c_int), ("dir_alt", c_int),
        ("mask_alt", c_void_p), ("mask_invert_alt", c_int),
        ("h1_out", c_void_p), ("h2_out", c_void_p),
    ]


class GemmSeg(Structure):
    """l2hmc_gemm_seg (include/l2hmc_b200.h)."""
    _fields_ = [("A", c_void_p), ("a_rs", c_longlong), ("a_cs", c_longlong), ("B", c_void_p), ("b_rs", c_longlong),
                ("b_cs", c_longlong), ("K", c_int)]


class GemmArgs(Structure):
    """l2hmc_gemm_args (include/l2hmc_b200.h)."""
    _fields_ = [("M", c_int), ("N", c_int), ("nseg", c_int), ("seg", GemmSeg * 3), ("C", c_void_p),
                ("ldc", c_longlong), ("accumulate", c_int), ("relu_gate", c_void_p), ("ldg", c_longlong),
                ("col_scale", c_void_p), ("col_scale_invert", c_int), ("row_set", c_void_p),
                ("col_scale_alt", c_void_p), ("col_scale_invert_alt", c_int)]


class UpdateBwdArgs(Structure):
    """l2hmc_update_bwd_args (include/l2hmc_b200.h)."""
    _fields_ = [("position", c_int), ("N", c_int), ("D", c_int), ("eps", c_float), ("dir", c_int), ("dir_alt", c_int),
                ("row_set", c_void_p), ("x", c_void_p), ("v", c_void_p), ("g", c_void_p),
                ("S", c_void_p), ("T", c_void_p), ("Q", c_void_p), ("coeff_s", c_void_p), ("coeff_q", c_void_p),
                ("mask", c_void_p), ("mask_alt", c_void_p), ("invert", c_int), ("invert_alt", c_int),
                ("gx", c_void_p), ("gv", c_void_p), ("gld", c_void_p),
                ("dA", c_void_p), ("dg", c_void_p), ("ps", c_void_p), ("pq", c_void_p), ("deps", c_void_p),
                ("eps_dev", c_void_p)]


class Conv3dWeights(Structure):
    """l2hmc_conv3d_weights (include/l2hmc_b200.h)."""
    _fields_ = [("w1", c_void_p), ("b1", c_void_p), ("w2", c_void_p), ("b2", c_void_p), ("F", c_int)]


class Conv3dTape(Structure):
    """l2hmc_conv3d_tape (include/l2hmc_b200.h)."""
    _fields_ = [("p1", c_void_p), ("arg1", c_void_p), ("arg2", c_void_p)]


class Conv3dBwdArgs(Structure):
    """l2hmc_conv3d_bwd_args (include/l2hmc_b200.h)."""
    _fields_ = [("w", POINTER(Conv3dWeights)), ("in_", c_void_p), ("mask", c_void_p), ("mask_invert", c_int),
                ("mask_alt", c_void_p), ("mask_invert_alt", c_int), ("row_set", c_void_p), ("N", c_int), ("T", c_int),
                ("X", c_int), ("g_feat", c_void_p), ("g_in", c_void_p), ("partial", c_void_p),
                ("tape", POINTER(Conv3dTape))]


class NetGrads(Structure):
    """l2hmc_net_grads (include/l2hmc_b200.h)."""
    _fields_ = [(n, c_void_p) for n in ("w_in0", "b_in0", "w_in1", "b_in1", "w_t", "b_t", "w_h", "b_h", "w_s", "b_s",
                                        "w_tr", "b_tr", "w_q", "b_q", "coeff_s", "coeff_q")]


class Conv3dGrads(Structure):
    """l2hmc_conv3d_grads (include/l2hmc_b200.h)."""
    _fields_ = [("w1", c_void_p), ("b1", c_void_p), ("w2", c_void_p), ("b2", c_void_p)]


class Gmm(Structure):
    """l2hmc_gmm (include/l2hmc_b200.h)."""
    _fields_ = [("k", c_int), ("m", c_int), ("mus", c_void_p), ("i_sigmas", c_void_p), ("log_consts", c_void_p)]


class L2hmcMcArgs(Structure):
    """l2hmc_l2hmc_mc_args (include/l2hmc_b200.h)."""
    _fields_ = [
        ("N", c_int), ("T", c_int), ("X", c_int), ("M", c_int), ("eps", c_float), ("beta", c_float),
        ("masks", c_void_p), ("ts", c_void_p), ("xnet", POINTER(NetWeights)), ("vnet", POINTER(NetWeights)),
        ("xconv_in0", POINTER(Conv3dWeights)), ("xconv_in1", POINTER(Conv3dWeights)),
        ("vconv_in0", POINTER(Conv3dWeights)), ("vconv_in1", POINTER(Conv3dWeights)), ("net_path", c_int),
        ("x", c_void_p), ("v_f", c_void_p), ("v_b", c_void_p), ("u_dir", c_void_p), ("u_acc", c_void_p),
        ("seed", c_uint64), ("mc_step", c_uint64), ("chain0", c_uint64), ("wrap_2pi", c_int),
        ("x_post", c_void_p), ("v_post", c_void_p), ("accept_prob", c_void_p), ("x_out", c_void_p),
        ("accepted", c_void_p), ("obs_action", c_void_p), ("obs_plaq", c_void_p), ("obs_charge", c_void_p),
        ("nonfinite_flag", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
        ("mc_step_dev", c_void_p),
    ]


class LossParams(Structure):
    """l2hmc_loss_params (include/l2hmc_b200.h)."""
    _fields_ = [("metric", c_int), ("loss_scale", c_float), ("aux_weight", c_float), ("std_weight", c_float),
                ("charge_weight", c_float)]


class TrainArgs(Structure):
    """l2hmc_train_args (include/l2hmc_b200.h)."""
    _fields_ = [
        ("N", c_int), ("T", c_int), ("X", c_int), ("M", c_int), ("eps", c_float), ("beta", c_float),
        ("masks", c_void_p), ("ts", c_void_p), ("xnet", POINTER(NetWeights)), ("vnet", POINTER(NetWeights)),
        ("xconv_in0", POINTER(Conv3dWeights)), ("xconv_in1", POINTER(Conv3dWeights)),
        ("vconv_in0", POINTER(Conv3dWeights)), ("vconv_in1", POINTER(Conv3dWeights)),
        ("x", c_void_p), ("z", c_void_p), ("v_f", c_void_p), ("v_b", c_void_p), ("u_dir", c_void_p), ("u_acc", c_void_p),
        ("loss", LossParams), ("forward_only", c_int),
        ("loss_out", c_void_p), ("per_chain", c_void_p), ("x_out", c_void_p), ("px", c_void_p), ("pz", c_void_p),
        ("x_prop", c_void_p), ("x_dq", c_void_p), ("accepted", c_void_p), ("g_eps", c_void_p),
        ("xnet_g", POINTER(NetGrads)), ("vnet_g", POINTER(NetGrads)),
        ("xconv_in0_g", POINTER(Conv3dGrads)), ("xconv_in1_g", POINTER(Conv3dGrads)),
        ("vconv_in0_g", POINTER(Conv3dGrads)), ("vconv_in1_g", POINTER(Conv3dGrads)),
        ("nonfinite_flag", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
        ("first_bucket_event", c_void_p), ("eps_dev", c_void_p),
    ]


class RunStats(Structure):
    """l2hmc_run_stats (include/l2hmc_b200.h)."""
    _fields_ = [("N", c_int), ("qmax", c_int), ("sum", c_void_p), ("sumsq", c_void_p), ("hist", c_void_p)]


METRICS = {"l1": 0, "l2": 1, "cos": 2, "cos2": 3, "cos_diff": 4}


class MogArgs(Structure):
    """l2hmc_mog_args (include/l2hmc_b200.h)."""
    _fields_ = [
        ("x", c_void_p), ("v_f", c_void_p), ("v_b", c_void_p), ("dir_mask", c_void_p), ("u_acc", c_void_p),
        ("mask", c_void_p), ("eps", c_float), ("temperature", c_float), ("L", c_int), ("hmc", c_int), ("N", c_int),
        ("seed", c_uint64), ("mc_step", c_uint64), ("chain0", c_uint64),
        ("xnet", POINTER(NetWeights)), ("vnet", POINTER(NetWeights)),
        ("Lx", c_void_p), ("Lv", c_void_p), ("px", c_void_p), ("x_out", c_void_p), ("accepted", c_void_p),
        ("mode", c_void_p), ("nonfinite_flag", c_void_p),
    ]


NET_STQ, NET_MOMENTUM, NET_POSITION = 0, 1, 2
NET_PATH_AUTO, NET_PATH_FFMA, NET_PATH_TCGEN05 = 0, 1, 2

# name -> argtypes; every function returns int status except the two noted below.
SIGNATURES = {
    "l2hmc_abi_version": [],
    "l2hmc_last_error": [],
    "l2hmc_launch_count": [],
    "l2hmc_device_info": [POINTER(c_int), POINTER(c_int)],
    "l2hmc_u1_action_force": [c_void_p, c_float, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p],
    "l2hmc_u1_observables": [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_u1_hamiltonian": [c_void_p, c_void_p, c_float, c_int, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_kinetic_energy": [c_void_p, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_update_momentum": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int, c_int, c_int,
                              c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_update_position": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float, c_int,
                              c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_mask_mul": [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_accept_prob": [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p],
    "l2hmc_mh_select": [c_void_p] * 9 + [c_int, c_int, c_int] + [c_void_p] * 5 + [c_void_p],
    "l2hmc_u1_hmc_step": [c_void_p, c_void_p, c_float, c_float, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                          c_void_p],
    "l2hmc_u1_hmc_mc_step": [POINTER(HmcMcArgs), c_void_p],
    "l2hmc_philox_draws": [c_uint64, c_uint64, c_uint64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_philox_draws_dev": [c_uint64, c_uint64, c_void_p, c_uint64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_counter_add": [c_void_p, c_uint64, c_void_p],
    "l2hmc_net_workspace_bytes": [c_int, c_int, c_int, c_int],
    "l2hmc_net_run": [POINTER(NetWeights), POINTER(NetCall), c_void_p],
    "l2hmc_net_pack_bytes": [c_int, c_int, c_int],
    "l2hmc_net_tc_min_rows": [c_int, c_int, c_int],
    "l2hmc_net_pack": [POINTER(NetWeights), c_void_p, c_size_t, c_void_p],
    "l2hmc_tc_selftest_bytes": [c_int, c_int, c_int],
    "l2hmc_tc_selftest": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p],
    "l2hmc_gmm_energy_force": [POINTER(Gmm), c_void_p, c_float, c_int, c_void_p, c_void_p, c_void_p],
    "l2hmc_mog_mc_step": [POINTER(Gmm), POINTER(MogArgs), c_void_p],
    "l2hmc_mog_loss_workspace_bytes": [],
    "l2hmc_mog_loss": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_longlong, c_int, c_float,
                       c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    "l2hmc_u1_l2hmc_mc_workspace_bytes": [POINTER(L2hmcMcArgs)],
    "l2hmc_u1_l2hmc_mc_step": [POINTER(L2hmcMcArgs), c_void_p],
    "l2hmc_u1_loss": [POINTER(LossParams), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                      c_int, c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_u1_top_charges_fft": [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_stats_update": [POINTER(RunStats), c_void_p, c_void_p, c_void_p, c_int, c_void_p],
    "l2hmc_stats_finalize": [POINTER(RunStats), c_longlong, c_longlong, c_void_p, c_void_p, c_void_p],
    "l2hmc_mode_switches": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p],
    "l2hmc_conv3d_num_features": [c_int, c_int, c_int],
    "l2hmc_conv3d_features": [POINTER(Conv3dWeights), c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                              c_void_p],
    "l2hmc_conv3d_tape_elems": [c_int, c_int, c_int],
    "l2hmc_conv3d_features_tape": [POINTER(Conv3dWeights), c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p,
                                   c_int, c_int, c_int, c_void_p, POINTER(Conv3dTape), c_void_p],
    "l2hmc_sgemm_workspace_bytes": [POINTER(GemmArgs), c_int],
    "l2hmc_sgemm": [POINTER(GemmArgs), c_int, c_void_p, c_size_t, c_void_p],
    "l2hmc_tc_gemm_workspace_bytes": [POINTER(GemmArgs)],
    "l2hmc_tc_gemm": [POINTER(GemmArgs), c_void_p, c_size_t, c_void_p],
    "l2hmc_gemm_workspace_bytes": [POINTER(GemmArgs), c_int],
    "l2hmc_gemm": [POINTER(GemmArgs), c_int, c_void_p, c_size_t, c_void_p],
    "l2hmc_train_update_fwd": [c_int] + [c_void_p] * 7 + [c_int, c_void_p, c_int, c_void_p, c_float, c_int, c_int,
                                                         c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_mask_mul_sets": [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_update_bwd": [POINTER(UpdateBwdArgs), c_void_p],
    "l2hmc_u1_force_hvp": [c_void_p, c_void_p, c_float, c_int, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_accept_prob_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p],
    "l2hmc_traj_bwd_init": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p],
    "l2hmc_u1_loss_bwd": [POINTER(LossParams), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                          c_void_p, c_void_p, c_void_p, c_void_p],
    "l2hmc_conv3d_bwd_rows": [c_int],
    "l2hmc_conv3d_bwd_cols": [c_int],
    "l2hmc_conv3d_backward": [POINTER(Conv3dBwdArgs), c_void_p],
    "l2hmc_u1_train_workspace_bytes": [POINTER(TrainArgs)],
    "l2hmc_u1_train_grads": [POINTER(TrainArgs), c_void_p],
    "l2hmc_col_sum_workspace_bytes": [c_int, c_int],
    "l2hmc_col_sum": [c_void_p, c_longlong, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p],
    "l2hmc_col_sum_static": [c_void_p, c_longlong, c_int, c_int, c_void_p, c_void_p],
    "l2hmc_reduce_sum": [c_void_p, c_longlong, c_int, c_void_p, c_void_p, c_void_p],
    "l2hmc_clip_by_global_norm": [c_void_p, c_longlong, c_void_p, c_float, c_void_p],
    "l2hmc_adam_step": [c_void_p, c_void_p, c_void_p, c_void_p, c_longlong, c_float, c_float, c_float, c_float,
                        c_float, c_void_p, c_float, c_void_p],
    "l2hmc_conv3d_features_sets": [POINTER(Conv3dWeights), c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p,
                                   c_int, c_int, c_int, c_void_p, c_void_p],
}

RESTYPES = {"l2hmc_last_error": c_char_p, "l2hmc_launch_count": ctypes.c_ulonglong, "l2hmc_net_workspace_bytes": c_size_t, "l2hmc_net_pack_bytes": c_size_t, "l2hmc_mog_loss_workspace_bytes": c_size_t,
            "l2hmc_tc_selftest_bytes": c_size_t, "l2hmc_u1_l2hmc_mc_workspace_bytes": c_size_t, "l2hmc_sgemm_workspace_bytes": c_size_t, "l2hmc_u1_train_workspace_bytes": c_size_t,
            "l2hmc_tc_gemm_workspace_bytes": c_size_t, "l2hmc_gemm_workspace_bytes": c_size_t, "l2hmc_col_sum_workspace_bytes": c_size_t}

L2HMC_OK, L2HMC_ERR_INVALID, L2HMC_ERR_UNSUPPORTED, L2HMC_ERR_CUDA = 0, 1, 2, 3
ABI_VERSION = 2          # L2HMC_ABI_VERSION of include/l2hmc_b200.h


class L2hmcError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"l2hmc_b200 error {code}: {text}")
        self.code = code


def load_library(path=LIB_PATH):
    """dlopen the library and bind every declared symbol (no device needed)."""
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python -m l2hmc_b200.build` (nvcc, sm_100a). "
            "l2hmc_b200 has no CPU or eager fallback.")
    lib = ctypes.CDLL(path)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = RESTYPES.get(name, c_int)
    return lib


def get_lib():
    global _lib
    if _lib is None:
        _lib = load_library()
        if _lib.l2hmc_abi_version() != ABI_VERSION:
            raise ImportError("libl2hmc_b200.so ABI version mismatch; rebuild")
    return _lib


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("l2hmc_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")


def check(rc):
    if rc != 0:
        raise L2hmcError(rc, get_lib().l2hmc_last_error().decode())


def stream_ptr():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def dptr(t, dtype=torch.float32):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous() and t.dtype == dtype):
        raise TypeError(f"expected a contiguous CUDA {dtype} tensor, got "
                        f"{type(t).__name__} {getattr(t, 'dtype', None)} {getattr(t, 'device', None)}")
    return c_void_p(t.data_ptr())


def launch_count():
    """Kernels launched by the library in this process so far (l2hmc_launch_count)."""
    return int(get_lib().l2hmc_launch_count())


def device_info():
    sms, smem = c_int(0), c_int(0)
    check(get_lib().l2hmc_device_info(ctypes.byref(sms), ctypes.byref(smem)))
    return sms.value, smem.value


FUSED_HMC_SHAPES = {(4, 4), (8, 8), (16, 16), (32, 32), (8, 16), (16, 32)}
