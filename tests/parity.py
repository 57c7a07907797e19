"""Shared parity helpers: the acceptance rule of SURVEY.md §8(c).

Per 8-bit channel |gpu - oracle| <= 1; pixels exceeding that must be <= 0.1 % of the frame."""
import numpy as np

MAX_LSB = 1          # tolerance per 8-bit channel (north_star: "within +-1 LSB")
MAX_OUTLIER_FRAC = 1e-3  # at most 0.1 % of pixels may differ more


def compare_frames(gpu_rgba: np.ndarray, oracle_rgba: np.ndarray) -> dict:
    assert gpu_rgba.shape == oracle_rgba.shape, (gpu_rgba.shape, oracle_rgba.shape)
    d = np.abs(gpu_rgba.astype(np.int16) - oracle_rgba.astype(np.int16)).max(axis=-1)
    n = d.size
    return {
        "pixels": int(n),
        "exact_frac": float((d == 0).sum()) / n,
        "within1_frac": float((d <= MAX_LSB).sum()) / n,
        "outliers": int((d > MAX_LSB).sum()),
        "outlier_frac": float((d > MAX_LSB).sum()) / n,
        "max_diff": int(d.max()) if n else 0,
    }


def assert_parity(gpu_rgba, oracle_rgba, what="", mask_fn=None):
    """The acceptance rule.  mask_fn (optional): a callable returning the frame's ill-conditioned-pixel mask, bool [H, W]
    (oracle_py.ill_conditioned_mask) — called only when some pixel exceeds +-1 LSB, and then every such pixel has to lie
    inside the mask ("only where the reference's own root solving is ill-conditioned", north_star)."""
    r = compare_frames(gpu_rgba, oracle_rgba)
    assert r["outlier_frac"] <= MAX_OUTLIER_FRAC, "{}: {}".format(what, r)
    assert (gpu_rgba[..., 3] == 255).all(), what + ": alpha"
    if r["outliers"] and mask_fn is not None:
        d = np.abs(gpu_rgba.astype(np.int16) - oracle_rgba.astype(np.int16)).max(axis=-1)
        outside = (d > MAX_LSB) & ~mask_fn()
        r["outliers_outside_mask"] = int(outside.sum())
        assert not outside.any(), "{}: {} outlier pixels outside the ill-conditioned mask, e.g. (y, x) = {}".format(
            what, int(outside.sum()), tuple(int(v) for v in np.argwhere(outside)[0]))
    return r
