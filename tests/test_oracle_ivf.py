"""oracle/ivf.py sanity on the CPU: probing every cluster equals a brute-force search over the reconstructions, and
the per-cluster bookkeeping partitions the ids (reference: src/haag_vq/methods/search/ivf_quantized_index.py:55-146)."""
import numpy as np

from oracle import ivf as oivf


class RoundQuantizer:
    """A deterministic toy quantizer (grid of step 1/8) with the BaseQuantizer surface."""
    def fit(self, X):
        self.D = X.shape[1]

    def compress(self, X):
        return np.clip(np.rint(X * 8.0), -127, 127).astype(np.int8).view(np.uint8)

    def decompress(self, codes):
        return codes.view(np.int8).astype(np.float32) / 8.0


def test_full_probe_equals_brute_force_and_ids_partition():
    rng = np.random.default_rng(0)
    X = rng.standard_normal((600, 16)).astype(np.float32)
    Q = rng.standard_normal((20, 16)).astype(np.float32)
    C = X[rng.choice(600, 12, replace=False)].copy()
    for metric in ("l2", "ip"):
        a, margin = oivf.assign(X, C, metric)
        assert margin.min() >= 0
        qs, codes, ids = oivf.build(X, C, a, RoundQuantizer)
        allids = np.sort(np.concatenate(ids))
        assert np.array_equal(allids, np.arange(600, dtype=np.uint32))
        got_ids, got_d, probe = oivf.search(Q, 7, C, qs, codes, ids, nprobe=12, metric=metric)
        recon = np.empty_like(X)
        for c in range(12):
            if qs[c] is not None:
                recon[ids[c]] = qs[c].decompress(codes[c]) + C[c]
        R64, Q64 = recon.astype(np.float64), Q.astype(np.float64)
        d = ((Q64[:, None, :] - R64[None]) ** 2).sum(2) if metric == "l2" else -(Q64 @ R64.T)
        ref = np.argsort(d, axis=1, kind="stable")[:, :7]
        assert np.array_equal(got_ids, ref.astype(np.uint32))
        assert np.allclose(got_d, np.take_along_axis(d, ref, 1))
        # fewer probes never return an id outside the probed clusters
        ids2, _, probe2 = oivf.search(Q, 7, C, qs, codes, ids, nprobe=3, metric=metric)
        for qi in range(Q.shape[0]):
            allowed = np.concatenate([ids[c] for c in probe2[qi]])
            assert np.isin(ids2[qi][ids2[qi] != np.iinfo(np.uint32).max], allowed).all()
