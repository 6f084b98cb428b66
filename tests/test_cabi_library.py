"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol the header
declares, and the host mirror refuses to run without a GPU (no silent fallback)."""
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "l2hmc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(l2hmc_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_loads_and_exports_every_declared_symbol():
    from l2hmc_b200 import _lib
    from l2hmc_b200.build import build
    lib = _lib.load_library(build())
    syms = header_symbols()
    assert len(syms) >= 14
    for name in syms:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(syms)
    assert lib.l2hmc_abi_version() == _lib.ABI_VERSION


def test_argument_validation_needs_no_device():
    from l2hmc_b200 import _lib
    lib = _lib.load_library()
    assert lib.l2hmc_u1_action_force(None, 1.0, 4, 8, 8, None, None, None) == _lib.L2HMC_ERR_INVALID
    assert b"null" in lib.l2hmc_last_error()
    assert lib.l2hmc_update_momentum(None, None, None, None, None, 0.1, 1, 1, 8, None, None, None, None) == 1


def test_missing_library_is_a_loud_import_error(tmp_path):
    from l2hmc_b200 import _lib
    with pytest.raises(ImportError, match="no CPU or eager fallback"):
        _lib.load_library(str(tmp_path / "nope.so"))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    import l2hmc_b200
    lat = l2hmc_b200.GaugeLattice(8, 8, 2, "U1", num_samples=2)     # host bookkeeping only
    assert lat.num_links == 128 and lat.num_plaquettes == 64 and lat.samples.shape == (2, 8, 8, 2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lat.total_action(np.zeros((2, 128), np.float32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        l2hmc_b200.GaugeDynamics(lat, hmc=True, num_steps=2, eps=0.1)


def test_host_helpers_match_oracle():
    from l2hmc_b200.dynamics import construct_masks, time_table
    from l2hmc_b200.lattice import u1_plaq_exact
    from oracle import u1_oracle as orc
    np.random.seed(42)
    m = construct_masks(128, 7)
    assert np.array_equal(m, orc.construct_masks(128, 7, seed=42)) and (m.sum(1) == 64).all()
    tt = time_table(5)
    for i in range(5):
        np.testing.assert_allclose(tt[i], orc.format_time(i, 5).numpy(), atol=1e-7)
    assert abs(u1_plaq_exact(2.0) - 0.6977746) < 1e-6


def test_dispatch_and_size_helpers_need_no_device():
    """Pure host functions of the library: the batch size from which the AUTO layer path takes tcgen05 (what the
    sampler's chunk rule is built on), the conv feature / tape sizes and the MoG loss workspace."""
    from l2hmc_b200 import _lib
    lib = _lib.get_lib()
    assert lib.l2hmc_net_tc_min_rows(128, 512, 128) == 1024          # 8x8 generic nets: max(256, 2^19 / 512)
    assert lib.l2hmc_net_tc_min_rows(512, 1024, 512) == 512          # 16x16 conv3D trunk
    assert lib.l2hmc_net_tc_min_rows(2048, 8192, 2048) == 256        # 32x32 generic nets
    assert lib.l2hmc_net_tc_min_rows(120, 480, 120) == 2 ** 31 - 1   # D % 16 != 0: never tcgen05
    assert lib.l2hmc_net_tc_min_rows(32, 512, 128) == 2 ** 31 - 1    # K < 64
    assert lib.l2hmc_conv3d_num_features(16, 16, 16) == 4 * 4 * 32
    assert lib.l2hmc_conv3d_tape_elems(16, 16, 16) == 8 * 8 * 16
    assert lib.l2hmc_mog_loss_workspace_bytes() >= 4 * 4 * 1024
