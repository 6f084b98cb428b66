"""l2hmc_b200: B200-native L2HMC / HMC leapfrog path for the 2D U(1) lattice gauge model.

Drop-in for the reference's GaugeLattice / GaugeDynamics / GenericNet / ConvNet3D surface
(saforem2/l2hmc); all arithmetic runs in hand-written sm_100a kernels behind the C ABI in
include/l2hmc_b200.h.  No CPU fallback: compute calls raise without the built library + a GPU.
"""
from .lattice import GaugeLattice, project_angle, u1_plaq_exact  # noqa: F401
from .dynamics import GaugeDynamics  # noqa: F401

__all__ = ["GaugeLattice", "GaugeDynamics", "u1_plaq_exact", "project_angle"]
