"""fvecs / ivecs I/O: byte-identical files to the reference's writers
(src/haag_vq/utils/faiss_export.py:46-133: int32 dimension prefix per record), round trips, and the
reference's error cases. CPU only: no kernel is involved."""
import numpy as np
import pytest

from haag_vq.utils.faiss_export import load_fvecs, load_ivecs, write_fvecs, write_ivecs


def _reference_bytes(vectors, dtype):
    # the reference's writer, restated: per row, int32(dim) then the row
    out = bytearray()
    for vec in np.asarray(vectors, dtype=dtype):
        out += np.int32(vec.shape[0]).tobytes() + np.asarray(vec, dtype=dtype).tobytes()
    return bytes(out)


@pytest.mark.parametrize("shape", [(1, 1), (5, 7), (33, 128), (0, 4)])
def test_fvecs_round_trip_and_bytes(tmp_path, shape):
    a = np.random.default_rng(1).standard_normal(shape).astype(np.float32)
    p = write_fvecs(tmp_path / "a.fvecs", a)
    assert p.read_bytes() == _reference_bytes(a, np.float32)
    back = load_fvecs(p)
    if shape[0] == 0:
        assert back.shape == (0, 0)      # empty file: the reference returns an empty (0, 0) matrix
    else:
        assert back.dtype == np.float32 and np.array_equal(back, a)


def test_ivecs_round_trip_and_bytes(tmp_path):
    b = np.random.default_rng(2).integers(-5, 1 << 20, size=(9, 3))
    p = write_ivecs(tmp_path / "b.ivecs", b)
    assert p.read_bytes() == _reference_bytes(b, np.int32)
    assert np.array_equal(load_ivecs(p), b.astype(np.int32))
    # non-contiguous input
    c = np.arange(24, dtype=np.float32).reshape(4, 6)[:, ::2]
    assert np.array_equal(load_fvecs(write_fvecs(tmp_path / "c.fvecs", c)), c)


def test_vec_io_errors(tmp_path):
    with pytest.raises(ValueError):
        write_fvecs(tmp_path / "x.fvecs", np.zeros(3, dtype=np.float32))
    with pytest.raises(FileNotFoundError):
        load_fvecs(tmp_path / "missing.fvecs")
    bad = tmp_path / "bad.fvecs"
    bad.write_bytes(np.array([3, 1, 2], dtype=np.int32).tobytes())     # truncated record
    with pytest.raises(ValueError):
        load_fvecs(bad)
    mixed = tmp_path / "mixed.ivecs"
    mixed.write_bytes(np.array([1, 7, 2, 8], dtype=np.int32).tobytes())  # dims 1 then 2
    with pytest.raises(ValueError):
        load_ivecs(mixed)
