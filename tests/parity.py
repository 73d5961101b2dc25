"""Image / hit-id comparison used by the parity tests (the north_star acceptance rule):
>= 99.9 % of pixels within 2/255 per channel, primary hit ids equal away from silhouettes."""
import numpy as np

TOL_LSB = 2           # per channel, of the R8G8B8A8_UNORM output
MAX_MISMATCH = 1e-3   # fraction of pixels allowed outside the tolerance


def mismatch_fraction(a_rgba8, b_rgba8, tol=TOL_LSB):
    d = np.abs(a_rgba8[..., :3].astype(np.int16) - b_rgba8[..., :3].astype(np.int16)).max(axis=-1)
    return float((d > tol).mean()), int(d.max())


def silhouette_mask(ids):
    """True where the 3x3 neighbourhood of the oracle id map holds more than one id."""
    h, w = ids.shape
    p = np.pad(ids, 1, mode="edge")
    m = np.zeros((h, w), bool)
    for dy in range(3):
        for dx in range(3):
            m |= p[dy:dy + h, dx:dx + w] != ids
    return m


def id_mismatches(oracle_ids, gpu_ids):
    """(mismatches away from silhouettes, mismatches overall)"""
    diff = oracle_ids != gpu_ids
    interior = ~silhouette_mask(oracle_ids)
    return int((diff & interior).sum()), int(diff.sum())


def report(oracle_img, gpu_img, oracle_ids=None, gpu_ids=None):
    frac, dmax = mismatch_fraction(oracle_img, gpu_img)
    frac0, _ = mismatch_fraction(oracle_img, gpu_img, tol=0)
    out = {"mismatch_frac_gt2": frac, "max_abs_diff": dmax, "frac_not_bit_equal": frac0}
    if oracle_ids is not None and gpu_ids is not None:
        a, b = id_mismatches(oracle_ids, gpu_ids)
        out["id_mismatch_interior"] = a
        out["id_mismatch_total"] = b
    return out
