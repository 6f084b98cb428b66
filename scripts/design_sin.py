"""Design + validate the branch-free fp32 sin used by the fused HMC kernels (csrc/fast_sin.cuh).
Emulates the fp32/FMA sequence in float64 (exact products) and reports max error vs float64 sin."""
import numpy as np
f32 = np.float32

def split_2pi():
    two_pi = 2 * np.pi
    c1 = f32(6.28125)                                   # 201/32: 8 significant bits -> j*c1 exact for |j| < 2^16
    r = two_pi - float(c1)
    c2 = f32(np.ldexp(np.round(np.ldexp(r, 20)), -20))  # ~11 significant bits
    c3 = f32(two_pi - float(c1) - float(c2))
    return c1, c2, c3

def fit_poly(deg_terms):
    # sin(r) = r + r*s*(S1 + S2 s + ...), s = r^2 on [0, pi/2]; weighted least squares on Chebyshev nodes
    n = 4000
    k = np.arange(n) + 0.5
    r = (np.pi / 2 + 1e-3) * (0.5 - 0.5 * np.cos(np.pi * k / n))
    r = r[r > 1e-4]
    s = r * r
    y = (np.sin(r) - r) / (r * s)
    A = np.stack([s ** i for i in range(deg_terms)], axis=1)
    coef = np.linalg.lstsq(A, y, rcond=None)[0]
    # a few Remez-like reweightings toward minimax of ABSOLUTE error of sin
    w = np.ones_like(r)
    for _ in range(30):
        W = w * (r * s)
        coef = np.linalg.lstsq(A * W[:, None], y * W, rcond=None)[0]
        err = np.abs((A @ coef - y) * r * s)
        w *= (1 + err / err.max())
    return [f32(c) for c in coef]

def fma(a, b, c):
    return f32(np.float64(a) * np.float64(b) + np.float64(c))

def fast_sin(x, C, S):
    c1, c2, c3 = C
    x = x.astype(f32)
    magic = f32(12582912.0)
    j = f32(fma(x, f32(1 / (2 * np.pi)), magic) - magic)
    r = fma(j, -c1, x); r = fma(j, -c2, r); r = fma(j, -c3, r)
    a = np.abs(r)
    pi_hi = f32(np.pi); pi_lo = f32(np.pi - float(pi_hi))
    folded = f32(f32(pi_hi - a) + pi_lo)
    a = np.where(a > f32(np.pi / 2), folded, a).astype(f32)
    r = np.where(np.signbit(r), -a, a).astype(f32)      # xor the sign bit (folded value may be < 0)
    s = f32(r * r)
    p = S[-1] * np.ones_like(s)
    for c in S[-2::-1]:
        p = fma(p, s, c)
    t = f32(r * s)
    return fma(t, p, r)

if __name__ == "__main__":
    C = split_2pi()
    print("C =", [repr(float(c)) for c in C], "sum err", float(C[0]) + float(C[1]) + float(C[2]) - 2 * np.pi)
    for terms in (4, 5):
        S = fit_poly(terms)
        rng = np.random.default_rng(0)
        for lo, hi in ((-np.pi, np.pi), (-50, 50), (-3000, 3000), (-40000, 40000)):
            x = rng.uniform(lo, hi, 2_000_000).astype(f32)
            y = fast_sin(x, C, S)
            ref = np.sin(x.astype(np.float64))
            err = np.abs(y.astype(np.float64) - ref)
            ulp = err / np.maximum(np.spacing(np.abs(ref).astype(f32)).astype(np.float64), 1e-45)
            print(f"terms={terms} range[{lo:.0f},{hi:.0f}] max abs err {err.max():.3e}  max ulp(of result) {ulp.max():.1f}  "
                  f"mean abs {err.mean():.2e}")
        print(" S =", [repr(float(c)) for c in S])


def fit_cos(deg_terms):
    n = 4000
    k = np.arange(n) + 0.5
    r = (np.pi / 2 + 1e-3) * (0.5 - 0.5 * np.cos(np.pi * k / n))
    r = r[r > 1e-3]
    s = r * r
    y = (np.cos(r) - 1.0) / s
    A = np.stack([s ** i for i in range(deg_terms)], axis=1)
    w = np.ones_like(r)
    for _ in range(30):
        W = w * s
        coef = np.linalg.lstsq(A * W[:, None], y * W, rcond=None)[0]
        err = np.abs((A @ coef - y) * s)
        w *= (1 + err / err.max())
    return [f32(c) for c in coef]


def fast_cos(x, C, Cc):
    c1, c2, c3 = C
    x = x.astype(f32)
    magic = f32(12582912.0)
    j = f32(fma(x, f32(1 / (2 * np.pi)), magic) - magic)
    r = fma(j, -c1, x); r = fma(j, -c2, r); r = fma(j, -c3, r)
    a = np.abs(r)
    pi_hi = f32(np.pi); pi_lo = f32(np.pi - float(pi_hi))
    folded = f32(f32(pi_hi - a) + pi_lo)
    fold = a > f32(np.pi / 2)
    a = np.where(fold, folded, a).astype(f32)
    s = f32(a * a)
    p = Cc[-1] * np.ones_like(s)
    for c in Cc[-2::-1]:
        p = fma(p, s, c)
    c = fma(s, p, f32(1.0))
    return np.where(fold, -c, c).astype(f32)


if __name__ == "__main__":
    for terms in (5, 6):
        Cc = fit_cos(terms)
        x = np.random.default_rng(1).uniform(-3000, 3000, 2_000_000).astype(f32)
        err = np.abs(fast_cos(x, split_2pi(), Cc).astype(np.float64) - np.cos(x.astype(np.float64)))
        print(f"cos terms={terms} max abs err {err.max():.3e} mean {err.mean():.2e}")
        print(" Cc =", [repr(float(c)) for c in Cc])
