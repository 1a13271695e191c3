"""Fits the minimax-ish polynomials used by the fast-math gather (atan on [0,1], tan on [0,pi/4]) and checks
them in emulated fp32 (Horner with fp32 rounding at every step).  Prints C initialisers."""
import numpy as np
from numpy.polynomial import chebyshev as Ch

f32 = np.float32

def fit(g, lo, hi, deg, n=4000):
    # Chebyshev nodes on [lo, hi]; least squares on relative error -> near-minimax after Remez-like reweighting
    k = np.arange(n)
    x = 0.5 * (lo + hi) + 0.5 * (hi - lo) * np.cos(np.pi * (k + 0.5) / n)
    y = g(x)
    w = 1.0 / np.abs(y)
    c = None
    for it in range(40):
        V = np.vander(x, deg + 1, increasing=True)
        c, *_ = np.linalg.lstsq(V * w[:, None], y * w, rcond=None)
        err = np.abs((V @ c - y) / y)
        w = w * (1 + 4 * err / err.max()) ** 0.5      # Lawson-style reweighting towards minimax
    return c

def horner32(c, s):
    r = np.full_like(s, f32(c[-1]), dtype=f32)
    for a in c[-2::-1]:
        r = (r.astype(np.float64) * s.astype(np.float64) + np.float64(f32(a))).astype(f32)   # fma in fp32
    return r

# atan(t) = t * P(t^2) on t in [0,1]
P = fit(lambda s: np.arctan(np.sqrt(s)) / np.sqrt(s), 1e-12, 1.0, 8)
t = np.linspace(1e-6, 1, 2000001).astype(f32)
s = (t.astype(np.float64) ** 2).astype(f32)
val = (horner32(P, s).astype(np.float64) * t.astype(np.float64)).astype(f32)
rel = np.abs(val.astype(np.float64) - np.arctan(t.astype(np.float64))) / np.arctan(t.astype(np.float64))
print("atan deg8 in s: max rel err fp32 =", rel.max())
print("static constexpr float ATAN_C[] = {" + ", ".join(f"{f32(a)!r}f".replace("np.float32(", "").replace(")", "") for a in P) + "};")

# tan(x) = x * Q(x^2) on [0, pi/4]
Q = fit(lambda s: np.tan(np.sqrt(s)) / np.sqrt(s), 1e-12, (np.pi / 4) ** 2 * 1.0001, 6)
x = np.linspace(1e-6, np.pi / 4, 2000001).astype(f32)
s = (x.astype(np.float64) ** 2).astype(f32)
val = (horner32(Q, s).astype(np.float64) * x.astype(np.float64)).astype(f32)
rel = np.abs(val.astype(np.float64) - np.tan(x.astype(np.float64))) / np.tan(x.astype(np.float64))
print("tan deg6 in s: max rel err fp32 =", rel.max())
print("static constexpr float TAN_C[] = {" + ", ".join(f"{f32(a)!r}f".replace("np.float32(", "").replace(")", "") for a in Q) + "};")
