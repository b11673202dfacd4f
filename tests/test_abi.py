"""The C-ABI library loads and exports every symbol include/b200vq.h declares (no GPU needed)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200vq.h")


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200vq_\w+)\s*\(", text)))


def test_library_built_in_tree():
    import b200vq
    assert os.path.exists(b200vq.LIB_PATH), "run __graft_entry__.build() first"
    assert os.path.dirname(b200vq.LIB_PATH).startswith(ROOT)  # in-tree, not site-packages


def test_every_declared_symbol_is_exported_and_bound():
    import b200vq
    lib = b200vq.load()
    syms = header_symbols()
    assert len(syms) >= 30
    nm = subprocess.run(["nm", "-D", "--defined-only", b200vq.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(b200vq_\w+)\b", nm))
    for s in syms:
        assert s in exported, f"{s} declared in b200vq.h but not exported by libb200vq.so"
        assert hasattr(lib, s)
        assert s in b200vq.SIGNATURES, f"{s} has no ctypes prototype"
    assert set(b200vq.SIGNATURES) == set(syms), "ctypes table and header disagree"


def test_no_torch_types_in_abi():
    text = open(HEADER).read()
    assert "torch" not in text.lower() and "at::" not in text
    assert 'extern "C"' in text


def test_abi_version_and_error_string():
    import b200vq
    lib = b200vq.load()
    assert lib.b200vq_abi_version() == 1
    assert isinstance(lib.b200vq_last_error(), bytes)


def test_sass_is_sm100a_only():
    import b200vq
    out = subprocess.run(["cuobjdump", "--list-elf", b200vq.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, f"unexpected architectures in the fat binary: {archs}"


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import numpy as np
    from b200vq import B200VQError
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    with pytest.raises(B200VQError):
        ScalarQuantizer(8).fit(np.zeros((4, 4), dtype=np.float32))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "vector-quantization_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), \
                    f"{f} imports the oracle: the product path must not route through it"
