"""Shared parity helpers for the GPU tests (the gates of BASELINE.json's north_star).

  * codes / ids: bit-exact except where the oracle's top-2 gap is below 1e-6 relative
  * MSE and estimated distances: within 1e-4 relative
"""
import numpy as np

GAP_REL = 1e-6     # north_star: exemption for near-ties
VAL_REL = 1e-4     # north_star: distances / MSE tolerance


def assert_codes_parity(got, ref, d1, d2, what="codes"):
    """got/ref (n, M) uint8; d1/d2 oracle best / second-best distance per (row, sub-quantizer)."""
    got, ref = np.asarray(got), np.asarray(ref)
    assert got.shape == ref.shape, f"{what}: shape {got.shape} vs {ref.shape}"
    diff = got != ref
    near_tie = (d2 - d1) <= GAP_REL * np.maximum(np.abs(d1), 1e-30)
    bad = diff & ~near_tie
    assert not bad.any(), (f"{what}: {int(bad.sum())} of {got.size} differ outside the 1e-6 top-2 exemption "
                           f"({int(diff.sum())} differ in total)")
    return int(diff.sum())


def assert_topk_parity(vals, ids, scores, k, metric, what="topk", gap_rel=GAP_REL, val_rel=VAL_REL, abs_tol=1e-7):
    """vals/ids (nq, k) from the device; scores (nq, n) oracle score matrix (fp32).
    metric 0 = L2 (ascending), 1 = IP (descending). Ids must equal the oracle's (value, id)-ordered
    top-k except at positions whose oracle value is within gap_rel of the competing value.
    abs_tol: absolute slack for scores formed by cancellation (|q|^2 + |x|^2 - 2 q.x near 0), given
    as eps * operand magnitude by the caller."""
    vals, ids = np.asarray(vals), np.asarray(ids).astype(np.int64)
    nq, n = scores.shape
    k = min(k, n)
    key = scores if metric == 0 else -scores
    n_mismatch = 0
    for q in range(nq):
        order = np.lexsort((np.arange(n), key[q]))[:k]
        ref_vals = scores[q][order]
        got_true = scores[q][ids[q]]          # oracle score of what the device returned
        scale = np.maximum(np.abs(ref_vals), 1e-30)
        # reported values agree with the oracle's value of the same id
        assert np.all(np.abs(vals[q] - got_true) <= val_rel * np.maximum(np.abs(got_true), 1e-6) + abs_tol), \
            f"{what}: query {q}: reported values deviate from the oracle by more than {val_rel}"
        mism = ids[q] != order
        if mism.any():
            n_mismatch += int(mism.sum())
            gap = np.abs(got_true - ref_vals)
            bad = mism & (gap > gap_rel * scale + abs_tol)
            assert not bad.any(), (f"{what}: query {q}: {int(bad.sum())} ids differ outside the {gap_rel} "
                                   f"tie exemption (positions {np.nonzero(bad)[0][:5]})")
        assert len(set(ids[q].tolist())) == k, f"{what}: query {q}: duplicate ids"
    return n_mismatch


def make_data(n, D, seed, normalize=True, aniso=True):
    """SURVEY.md §8(d) synthetic recipe: anisotropic Gaussian, optionally row-normalised."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, D)).astype(np.float32)
    if aniso:
        X *= np.linspace(4.0, 0.2, D, dtype=np.float32)
        A = np.random.default_rng(12345).standard_normal((D, D)).astype(np.float32)
        X = (X @ A) / np.float32(np.sqrt(D))
    if normalize:
        X /= np.maximum(np.linalg.norm(X, axis=1, keepdims=True), 1e-12)
    return np.ascontiguousarray(X, dtype=np.float32)
