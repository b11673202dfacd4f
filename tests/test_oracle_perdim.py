"""oracle/perdim.py and the host training logic of haag_vq's RankAwareQuantizer (Lloyd-Max tables, greedy
allocation) against the reference's golden state (oracle/make_golden_perdim.py; reference
src/haag_vq/methods/rank_aware_quantization.py). CPU only."""
import os

import numpy as np

from oracle import perdim as opd


def _state(g):
    cb, o = [], 0
    for n in g["cb_len"]:
        cb.append(g["cb_flat"][o:o + n])
        o += n
    return g["mu"], g["V"], g["bits"], cb


def test_oracle_reproduces_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "perdim_D24.npz"))
    mu, V, bits, cb = _state(g)
    for packing in ("dense", "ffd"):
        codes = opd.compress(g["X"], mu, V, bits, cb, packing)
        assert codes.shape[1] == int(g[f"code_size_{packing}"])
        assert np.array_equal(codes, g[f"codes_{packing}"])
        assert np.array_equal(opd.decompress(codes, mu, V, bits, cb, packing), g[f"rec_{packing}"])
    assert np.array_equal(g["rec_dense"], g["rec_ffd"])      # FFD is a lossless re-layout (reference test_ffd_roundtrip)


def test_host_training_logic_matches_reference(golden_dir):
    from haag_vq.methods.rank_aware_quantization import RankAwareQuantizer, _lloyd_1d_normal
    g = np.load(os.path.join(golden_dir, "perdim_D24.npz"))
    Dg = np.empty(9)
    for b in range(9):
        lv, Dg[b] = _lloyd_1d_normal(2 ** b, seed=b)
        assert np.array_equal(lv, g[f"levels{b}"])
    Dg[0] = 1.0
    assert np.array_equal(Dg, g["Dg"])
    q = RankAwareQuantizer(avg_bits=3.5, alpha=0.5)
    bits = q._allocate(g["var"], Dg, int(round(3.5 * 24)))
    assert np.array_equal(bits, g["bits"]) and bits.sum() <= 84
