"""Fit the odd minimax polynomial used by atan2_(float, float) in csrc/salp_step.cuh:
atan(a) ~= a + a^3 * P(a^2) on [0, 1] (Lawson-weighted least squares), and report the max abs error
when evaluated in float32 Horner form.  Developer tool; prints the coefficients (lowest order first)."""
import numpy as np

a = np.unique(np.concatenate([np.cos(np.linspace(0, np.pi / 2, 20001)), np.linspace(0, 1, 20001)]))
a = a[a > 1e-6]
s = a * a
target = (np.arctan(a) - a) / a ** 3
for deg in (4, 5, 6, 7):
    V = np.vander(s, deg + 1, increasing=True)
    w = np.ones_like(a)
    for _ in range(200):
        c, *_ = np.linalg.lstsq(V * (w * a ** 3)[:, None], target * w * a ** 3, rcond=None)
        err = np.abs((V @ c) * a ** 3 + a - np.arctan(a))
        w = w * (0.5 + err / err.max())
        w /= w.mean()
    c32 = c.astype(np.float32)
    x = np.linspace(0, 1, 2000001).astype(np.float32)
    ss = x * x
    r = np.full_like(x, c32[-1])
    for k in range(deg - 1, -1, -1):
        r = r * ss + c32[k]
    r = (r * ss) * x + x
    e32 = np.abs(r.astype(np.float64) - np.arctan(x.astype(np.float64))).max()
    print(deg, f"f64 max err {err.max():.3e}  f32-evaluated max err {e32:.3e}", [float(v) for v in c32])
