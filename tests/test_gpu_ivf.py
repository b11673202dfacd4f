"""IVF shell on the GPU (haag_vq.methods.search.ivf_quantized_index, csrc/ivf.cu) against oracle/ivf.py, the numpy
restatement of src/haag_vq/methods/search/ivf_quantized_index.py:55-146. The coarse centroids (faiss.Kmeans in the
reference, restated on the device) are handed to the oracle; everything computed from them must agree: assignments
(except numerical ties), per-cluster codes, and the search results (ids identical except ties within 1e-6,
distances within 1e-4 relative)."""
import os
import pickle

import numpy as np
import pytest

from oracle import ivf as oivf

pytestmark = pytest.mark.gpu


class RoundQuantizer:
    """Deterministic toy quantizer (grid of step 1/16): the same object on both sides isolates the IVF logic."""
    def fit(self, X):
        self.D = X.shape[1]

    def compress(self, X):
        return np.clip(np.rint(X * 16.0), -127, 127).astype(np.int8).view(np.uint8)

    def decompress(self, codes):
        return codes.view(np.int8).astype(np.float32) / 16.0


def _data(n, D, nq, seed):
    rng = np.random.default_rng(seed)
    centers = rng.standard_normal((24, D)).astype(np.float32) * 2.0
    X = (centers[rng.integers(0, 24, n)] + 0.6 * rng.standard_normal((n, D))).astype(np.float32)
    Q = (centers[rng.integers(0, 24, nq)] + 0.6 * rng.standard_normal((nq, D))).astype(np.float32)
    return X, Q


def _check_search(index, X, Q, k, factory, metric, nprobe):
    C = index._centroids
    a_ref, margin = oivf.assign(X, C, metric)
    a_got = np.empty(X.shape[0], dtype=np.int32)
    for c, vid in enumerate(index._cluster_ids):
        a_got[vid] = c
    scale = np.abs(X).max() ** 2 * X.shape[1]
    diff = a_got != a_ref
    assert (margin[diff] <= 1e-5 * scale).all(), "assignment differs beyond numerical ties"
    # the oracle continues from the device's assignment so that one tie does not cascade
    qs, codes, ids = oivf.build(X, C, a_got, factory)
    for c in range(len(ids)):
        assert np.array_equal(ids[c], index._cluster_ids[c])
        assert np.array_equal(np.asarray(codes[c]), np.asarray(index._cluster_codes[c])), f"codes of cluster {c}"
    ref_ids, ref_d, _ = oivf.search(Q, k, C, qs, codes, ids, nprobe, metric)
    got_ids, got_d = index.search_with_scores(Q, k)
    assert got_ids.dtype == np.uint32 and got_d.dtype == np.float32 and got_ids.shape == (Q.shape[0], k)
    missing = ref_ids == np.iinfo(np.uint32).max
    assert np.array_equal(got_ids == np.iinfo(np.uint32).max, missing)
    fmax = np.finfo(np.float32).max
    assert (got_d[missing] == (-fmax if metric == "ip" else fmax)).all()
    ok = ~missing
    tol = 1e-4 * max(1.0, float(np.abs(ref_d[ok]).max()))
    assert np.abs(got_d[ok] - ref_d[ok]) .max() <= tol
    bad = (got_ids != ref_ids) & ok
    if bad.any():
        # only exchanges between (near-)equal distances are allowed: the probed-cluster boundary and fp32 rounding
        qi, j = np.nonzero(bad)
        for a, b in zip(qi, j):
            nb = ref_d[a, max(0, b - 1):b + 2]
            assert np.abs(nb - ref_d[a, b]).min() <= tol or np.abs(got_d[a, b] - ref_d[a, b]) <= 1e-6 * max(1.0, abs(ref_d[a, b]))
        assert bad.mean() < 0.01


@pytest.mark.parametrize("metric", ["l2", "ip"])
def test_ivf_matches_oracle_with_toy_quantizer(metric):
    from haag_vq.methods.search.ivf_quantized_index import IvfQuantizedIndex
    X, Q = _data(6000, 48, 70, 1)
    idx = IvfQuantizedIndex(RoundQuantizer, K=40, nprobe=6)
    idx.fit(X, metric)
    assert idx._centroids.shape == (40, 48)
    sizes = np.array([v.size for v in idx._cluster_ids])
    assert sizes.sum() == 6000 and (sizes > 0).sum() >= 30
    _check_search(idx, X, Q, 10, RoundQuantizer, metric, 6)
    assert idx.memory_footprint() == idx._centroids.nbytes + sum(c.nbytes for c in idx._cluster_codes)
    mse = idx.reconstruction_mse(X, np.arange(0, 6000, 7))
    assert mse is not None and 0 < mse < 1e-2


def test_ivf_with_device_quantizers_given_centroids_empty_clusters_and_save_load(tmp_path):
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    from haag_vq.methods.search.ivf_quantized_index import IvfQuantizedIndex
    X, Q = _data(4000, 64, 40, 2)
    rng = np.random.default_rng(3)
    C = X[rng.choice(4000, 20, replace=False)].copy()
    C[7] = 1000.0                                  # a centroid nobody is assigned to: empty cluster
    factory = lambda: ScalarQuantizer(num_bits=8)
    idx = IvfQuantizedIndex(factory, K=20, nprobe=20)     # every cluster probed: includes the empty one
    idx.fit(X, "l2", centroids=C)
    assert idx._cluster_quantizers[7] is None and idx._cluster_ids[7].size == 0
    _check_search(idx, X, Q, 100, factory, "l2", 20)
    ids0, d0 = idx.search_with_scores(Q, 5)
    # k larger than the probed rows: sentinels
    small = IvfQuantizedIndex(factory, K=20, nprobe=1)
    small.fit(X[:60], "l2", centroids=C)
    ids_s, d_s = small.search_with_scores(Q[:5], 128)
    assert (ids_s == np.iinfo(np.uint32).max).any() and (d_s[ids_s == np.iinfo(np.uint32).max] == np.finfo(np.float32).max).all()
    # any k: pages of 128 chained on the device equal the oracle's loop
    _check_search(idx, X, Q, 300, factory, "l2", 20)
    assert idx.search_with_scores(Q, 0)[0].shape == (40, 0)
    # save -> load -> identical results (the reference's pickle layout)
    path = os.path.join(tmp_path, "ivf.pkl")
    idx.save(path)
    st = pickle.load(open(path, "rb"))
    assert set(st) == {"K", "nprobe", "centroids", "serialised_quantizers", "cluster_codes", "cluster_ids", "metric",
                       "N", "D"}
    assert st["serialised_quantizers"][7] == {"type": "none"}
    other = IvfQuantizedIndex(factory, K=1, nprobe=1)
    other.load(path)
    ids1, d1 = other.search_with_scores(Q, 5)
    assert np.array_equal(ids0, ids1) and np.array_equal(d0, d1)


def test_ivf_nprobe_above_128_pq_clusters_and_recall():
    """nprobe > 128 exercises the chained exact probe (b200vq_flat_topk_after); PQ per cluster; recall vs exact."""
    from haag_vq.methods.product_quantization import ProductQuantizer
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    from haag_vq.methods.search.ivf_quantized_index import IvfQuantizedIndex
    from b200vq import host, ops
    X, Q = _data(60000, 32, 100, 4)
    _, gt = ops.flat_topk(host.to_device(X), host.to_device(Q), 10, "l2")
    gt = host.to_host(gt).view(np.uint32)
    idx = IvfQuantizedIndex(lambda: ScalarQuantizer(num_bits=8), K=200, nprobe=150)
    idx.fit(X, "l2")
    # the probe lists are the exact ranking of the centroids
    probe = idx._nearest_centroids(Q)
    assert probe.shape == (100, 150)
    C64 = idx._centroids.astype(np.float64)
    cs = ((Q.astype(np.float64)[:, None, :] - C64[None]) ** 2).sum(2)
    ref = np.argsort(cs, axis=1, kind="stable")[:, :150]
    assert (probe == ref).mean() > 0.995
    assert all(len(set(p)) == 150 for p in probe)
    ids, d = idx.search_with_scores(Q, 10)
    recall = np.mean([len(set(ids[i]) & set(gt[i])) / 10 for i in range(100)])
    assert recall > 0.9, recall
    assert (np.diff(d, axis=1) >= 0).all()
    pq = IvfQuantizedIndex(lambda: ProductQuantizer(M=8, B=8), K=16, nprobe=4)
    pq.fit(X, "l2")
    ids, d = pq.search_with_scores(Q, 10)
    recall = np.mean([len(set(ids[i]) & set(gt[i])) / 10 for i in range(100)])
    assert recall > 0.3, recall


@pytest.mark.parametrize("metric", ["l2", "ip"])
def test_ivf_tensor_core_search_equals_the_simt_kernel(metric):
    """b200vq_ivf_search_tc (screen + candidate pool with the probe lists as a mask) == b200vq_ivf_search, bit for bit."""
    import torch
    from b200vq import host, ivf, ops
    X, Q = _data(70000, 64, 300, 7)
    rng = np.random.default_rng(8)
    K, nprobe = 64, 8
    C = X[rng.choice(70000, K, replace=False)].copy()
    xd, qd, cd = host.to_device(X), host.to_device(Q), host.to_device(C)
    a = ivf.CoarseAssigner(cd, metric == "ip")(xd).long()
    order = torch.argsort(a, stable=True)
    recon = xd[order].contiguous()                    # "reconstructions" = the rows themselves
    rowc = a[order].int().contiguous()
    gid = order.int().contiguous()
    row_off = torch.zeros(K + 1, dtype=torch.int64, device=xd.device)
    row_off[1:] = torch.cumsum(torch.bincount(a, minlength=K), 0)
    rnorm = ops.row_sqnorms(recon)
    _, probe = ops.flat_topk(cd, qd, nprobe, metric, impl="simt")
    for k in (1, 10, 100):
        v0, i0 = ivf.ivf_search(recon, rnorm, gid, row_off, K, qd, probe, k, metric, impl="simt")
        v1, i1 = ivf.ivf_search(recon, rnorm, gid, row_off, K, qd, probe, k, metric, rowcluster=rowc, impl="tc")
        assert torch.equal(i0, i1), (metric, k, (i0 != i1).sum().item())
        assert torch.equal(v0.view(torch.int32), v1.view(torch.int32))
    # every returned row belongs to a probed cluster
    ids = i1.cpu().numpy().view(np.uint32)
    assign = a.cpu().numpy()
    pr = probe.cpu().numpy()
    for q in range(0, 300, 17):
        got = ids[q][ids[q] != np.iinfo(np.uint32).max]
        assert np.isin(assign[got], pr[q]).all()


@pytest.mark.parametrize("metric", ["l2", "ip"])
def test_ivf_any_k_by_chained_pages(metric):
    """k > 128 = pages chained by b200vq_ivf_search_after: equal to the oracle, continuing the (distance, id) order
    across identical rows, and sentinels once the probed rows run out."""
    from haag_vq.methods.search.ivf_quantized_index import IvfQuantizedIndex
    X, Q = _data(5000, 32, 33, 11)
    X[1000:1200] = X[999]                             # 201 identical rows straddle a page boundary
    idx = IvfQuantizedIndex(RoundQuantizer, K=12, nprobe=3)
    idx.fit(X, metric)
    _check_search(idx, X, Q, 400, RoundQuantizer, metric, 3)
    ids, d = idx.search_with_scores(Q, 400)
    for r in range(Q.shape[0]):
        valid = ids[r] != np.iinfo(np.uint32).max
        assert len(set(ids[r][valid])) == valid.sum()
        assert np.array_equal(np.lexsort((ids[r][valid], d[r][valid])), np.arange(valid.sum()))
    tiny = IvfQuantizedIndex(RoundQuantizer, K=4, nprobe=1)
    tiny.fit(X[:300], metric)
    ids, d = tiny.search_with_scores(Q[:4], 290)
    assert (ids == np.iinfo(np.uint32).max).any()
