"""Stated tolerances (BASELINE.json north_star: "matching the reference within 1e-4 relative
magnitude"; integer tables, bin indices bit-exact).

Two float32 FFTs of the same input differ by O(eps * log2 N) of the frame's largest bin, not of
each bin, so a purely relative bound cannot hold on bins 100 dB under the peak.  Linear magnitudes
are therefore compared with  |a - b| <= REL * |b| + FLOOR * max|b|  per frame:
1e-4 relative with an absolute floor 120 dB under the frame peak.  Log-scaled results live on the
reference's [0, 1] display scale (1.0 = highDBs, 0.0 = lowDBs) and are compared absolutely."""
import numpy as np

REL = 1e-4
FLOOR = 1e-6
LOG_ABS = 1e-4


def assert_linear_close(a, b, rel=REL, floor=FLOOR, frame_axes=None, what=""):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    peak = np.max(np.abs(b), axis=frame_axes, keepdims=True) if frame_axes is not None else np.max(np.abs(b))
    bound = rel * np.abs(b) + floor * peak
    err = np.abs(a - b)
    bad = err > bound
    if np.any(bad):
        i = np.unravel_index(np.argmax(err - bound), err.shape)
        raise AssertionError(f"{what}: {bad.sum()} of {bad.size} outside tolerance; worst at {i}: got {a[i]!r} want {b[i]!r} "
                             f"err {err[i]:.3e} bound {np.broadcast_to(bound, err.shape)[i]:.3e}")


def assert_log_close(a, b, lin=None, peak=None, dy=1.0 / np.log(1e6), rel=REL, floor=FLOOR, what=""):
    """Results on the display scale y = ln(v / low) * dy.  The linear bound |dv| <= x v with x = rel + floor * peak / v
    maps to |dy| <= -dy ln(1 - x), exactly (no bound where x >= 1: the value may be anything down to silence);
    2e-6 is added for the float32 rounding of y itself.  Without `lin` a flat LOG_ABS is used."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    err = np.abs(a - b)
    if lin is None:
        bound = np.full(err.shape, LOG_ABS)
    else:
        lin = np.asarray(lin, dtype=np.float64)
        pk = np.max(lin) if peak is None else peak
        with np.errstate(divide="ignore"):
            ratio = np.where(lin > 0, floor * pk / np.where(lin > 0, lin, 1.0), np.inf)
        x = rel + ratio
        with np.errstate(divide="ignore", invalid="ignore"):
            bound = np.where(x < 1.0, -dy * np.log1p(-np.minimum(x, 0.999999)), np.inf) + 2e-6
    bad = err > bound
    if np.any(bad):
        i = np.unravel_index(np.argmax(err - np.where(np.isfinite(bound), bound, 0)), err.shape)
        raise AssertionError(f"{what}: {bad.sum()} of {bad.size} outside tolerance; worst at {i}: got {a[i]!r} want {b[i]!r} bound {bound[i]:.2e}")
