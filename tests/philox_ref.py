"""NumPy restatement of the Philox4x32-10 + Box-Muller stream the kernels use
(csrc/common.cuh: Philox, philox_chain, box_muller).  Test infrastructure."""
import numpy as np

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF


def philox4x32_10(c, k):
    """c: (..., 4) uint64 holding 32-bit words; k: (2,) ints."""
    c0, c1, c2, c3 = (c[..., i].astype(np.uint64) for i in range(4))
    k0, k1 = k
    for _ in range(10):
        p0 = np.uint64(M0) * c0
        p1 = np.uint64(M1) * c2
        n0 = (p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)
        n1 = p1 & np.uint64(MASK)
        n2 = (p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)
        n3 = p0 & np.uint64(MASK)
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return np.stack([c0, c1, c2, c3], axis=-1)


def _counters(seed, mc_step, chains, idx):
    chains = np.asarray(chains, dtype=np.uint64)
    c = np.zeros(chains.shape + (4,), dtype=np.uint64)
    c[..., 0] = idx
    c[..., 1] = chains & np.uint64(MASK)
    c[..., 2] = np.uint64(mc_step & MASK)
    c[..., 3] = ((chains >> np.uint64(32)) << np.uint64(16)) | np.uint64((mc_step >> 32) & 0xFFFF)
    return c, (seed & MASK, (seed >> 32) & MASK)


def philox_uniforms(seed, mc_step, chain0, N):
    c, k = _counters(seed, mc_step, chain0 + np.arange(N), 0xFFFFFFFF)
    r = philox4x32_10(c, k)
    u = (r >> np.uint64(8)).astype(np.float32) * np.float32(2.0 ** -24)
    return u[:, 0], u[:, 1]


def philox_normals(seed, mc_step, chain0, N, D):
    quads = (D + 3) // 4
    chains = np.repeat(chain0 + np.arange(N), quads)
    idx = np.tile(np.arange(quads), N)
    c, k = _counters(seed, mc_step, chains, idx.astype(np.uint64))
    r = philox4x32_10(c, k)
    out = np.empty((N * quads, 4), dtype=np.float64)
    for j in (0, 2):
        u1 = ((r[:, j] >> np.uint64(8)).astype(np.float64) + 1.0) * 2.0 ** -24
        u2 = (r[:, j + 1] >> np.uint64(8)).astype(np.float64) * 2.0 ** -24
        u1 = u1.astype(np.float32).astype(np.float64)
        rad = np.sqrt(-2.0 * np.log(u1))
        ang = (np.float64(np.float32(2 * np.pi)) * u2).astype(np.float32)       # fp32 product,
        ang = (ang - np.float32(np.pi)).astype(np.float32).astype(np.float64)    # then fp32 shift to [-pi, pi)
        out[:, j] = rad * np.cos(ang)
        out[:, j + 1] = rad * np.sin(ang)
    return out.reshape(N, quads * 4)[:, :D].astype(np.float32)
