"""Row-shard exchange logic (b200vq/dist.py) on CPU: world size 2, gloo backend.

The data-path kernels need a GPU; what is covered here is the N > 1 plumbing: shard bounds,
local -> global ids, the all-gather of per-shard top-k lists, the (value, id) merge order and
the all-reduce of training statistics. The merge itself is the numpy twin of the device
kernel's ordering (dist.numpy_merge); the oracle supplies the expected global top-k."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as tdist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, metric, out_dir):
    for p in (os.path.join(ROOT, "vector-quantization_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    tdist.init_process_group("gloo", rank=rank, world_size=world)
    from b200vq import dist as vdist
    from oracle import pq as opq

    n_total, D, M, nq, k = 1001, 32, 4, 7, 10
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n_total, D)).astype(np.float32)
    Q = rng.standard_normal((nq, D)).astype(np.float32)
    C = opq.pq_init_centroids(X, M, 256, seed=3)
    codes = opq.pq_assign(X, C)
    codes[17] = codes[900]  # identical codes in two different shards: the id must break the tie
    s, e = vdist.shard_bounds(n_total, world, rank)
    lut = opq.adc_lut(Q, C, metric)
    sc = opq.adc_scores(codes[s:e], lut)
    v, i = opq.topk_sorted(sc, k, metric)
    vals, ids = vdist.sharded_topk(torch.from_numpy(v), torch.from_numpy(i.astype(np.int64).astype(np.int32)),
                                   s, k, metric, merge_fn=vdist.numpy_merge)
    # training statistics
    sums = torch.full((2, 3), float(rank + 1))
    counts = torch.full((2,), rank + 1, dtype=torch.int32)
    vdist.allreduce_kmeans_stats(sums, counts)
    vmin, vmax = torch.tensor([float(rank)]), torch.tensor([float(rank)])
    vdist.allreduce_minmax(vmin, vmax)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), vals=vals.numpy(), ids=ids.numpy(), sums=sums.numpy(),
             counts=counts.numpy(), vmin=vmin.numpy(), vmax=vmax.numpy())
    tdist.barrier()
    tdist.destroy_process_group()


@pytest.mark.parametrize("metric", [0, 1])
def test_sharded_topk_world2_matches_unsharded_oracle(tmp_path, metric):
    from oracle import pq as opq
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), metric, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (np.load(tmp_path / f"r{r}.npz") for r in range(world))
    assert np.array_equal(r0["ids"], r1["ids"]) and np.array_equal(r0["vals"], r1["vals"])
    n_total, D, M, nq, k = 1001, 32, 4, 7, 10
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n_total, D)).astype(np.float32)
    Q = rng.standard_normal((nq, D)).astype(np.float32)
    C = opq.pq_init_centroids(X, M, 256, seed=3)
    codes = opq.pq_assign(X, C)
    codes[17] = codes[900]
    v, i = opq.topk_sorted(opq.adc_scores(codes, opq.adc_lut(Q, C, metric)), k, metric)
    assert np.array_equal(r0["ids"], i.astype(np.int64))
    assert np.array_equal(r0["vals"], v)
    assert np.all(r0["sums"] == 3.0) and np.all(r0["counts"] == 3)
    assert r0["vmin"][0] == 0.0 and r0["vmax"][0] == 1.0


def test_shard_bounds_cover_rows_exactly():
    sys.path.insert(0, os.path.join(ROOT, "vector-quantization_b200"))
    from b200vq.dist import shard_bounds
    for n, w in ((53_000_000, 8), (1001, 2), (7, 8), (0, 4)):
        b = [shard_bounds(n, w, r) for r in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[r][1] == b[r + 1][0] for r in range(w - 1))
        assert max(e - s for s, e in b) - min(e - s for s, e in b) <= 1
