"""GPU parity tests: scalar quantizer (bit-exact vs the reference's golden vectors) and SAQ."""
import os

import numpy as np
import pytest

from oracle import sq as osq
from parity import make_data

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import ops_sq
    b200vq.load()
    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops_sq = torch, ops_sq
    ns.dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return ns


@pytest.mark.parametrize("name", ["sq_D48.npz", "sq_D33.npz"])
def test_sq_bit_exact_vs_reference_golden(T, golden_dir, name):
    g = np.load(os.path.join(golden_dir, name))
    vmin, vmax = T.ops_sq.sq_minmax(T.dev(g["X"]))
    assert np.array_equal(vmin.cpu().numpy(), g["vmin"]) and np.array_equal(vmax.cpu().numpy(), g["vmax"])
    D = g["X"].shape[1]
    for bits in (4, 8, 16):
        codes = T.ops_sq.sq_encode(T.dev(g["Xq"]), vmin, vmax, bits)
        got = codes.cpu().numpy()
        if bits == 16:
            got = got.view(np.uint16)
        assert np.array_equal(got, g[f"codes{bits}"]), f"SQ{bits} codes"
        rec = T.ops_sq.sq_decode(codes, vmin, vmax, bits, D).cpu().numpy()
        assert np.array_equal(rec, g[f"recon{bits}"]), f"SQ{bits} reconstruction"


@pytest.mark.parametrize("n,D", [(20000, 3072), (1001, 100), (5, 7)])
def test_sq_bit_exact_vs_oracle_at_size(T, n, D):
    X = make_data(n, D, seed=D, normalize=False, aniso=False) * np.linspace(0.1, 5, D, dtype=np.float32)
    vmin, vmax = osq.sq_fit(X)
    xd = T.dev(X)
    gmin, gmax = T.ops_sq.sq_minmax(xd)
    assert np.array_equal(gmin.cpu().numpy(), vmin) and np.array_equal(gmax.cpu().numpy(), vmax)
    for bits in (4, 8, 16):
        ref = osq.sq_compress(X, vmin, vmax, bits)
        codes = T.ops_sq.sq_encode(xd, gmin, gmax, bits)
        got = codes.cpu().numpy()
        got = got.view(np.uint16) if bits == 16 else got
        assert np.array_equal(got, ref)
        rec = T.ops_sq.sq_decode(codes, gmin, gmax, bits, D).cpu().numpy()
        assert np.array_equal(rec, osq.sq_decompress(ref, vmin, vmax, bits))


@pytest.mark.parametrize("bits", [4, 8, 16])
def test_sq_constant_division_is_ieee_division(T, bits):
    """The vector kernels divide by a per-column constant with reciprocal + one correction step
    (csrc/sq.cu div_const); the codes must be those of IEEE division on every input: 2^26 random
    (x - min, range) pairs with ranges from 1e-7 to 1e7, level boundaries, out-of-range and special values."""
    torch = T.torch
    g = torch.Generator(device="cuda").manual_seed(100 + bits)
    n, D = 1 << 14, 4096
    lm1 = float((1 << bits) - 1)
    vmin = (torch.rand(D, generator=g, device="cuda") - 0.5) * torch.pow(10.0, torch.rand(D, generator=g, device="cuda") * 8 - 4)
    span = torch.pow(10.0, torch.rand(D, generator=g, device="cuda") * 14 - 7)
    vmax = vmin + span
    u = torch.rand((n, D), generator=g, device="cuda") * 1.2 - 0.1
    # a third of the rows sit on level boundaries (k + 0.5) / (L - 1), where rounding decides the code
    kk = torch.randint(0, int(lm1), (n // 3, D), generator=g, device="cuda").float()
    u[: n // 3] = (kk + 0.5) / lm1
    x = vmin + u * (vmax - vmin)
    x[-1, :8] = torch.tensor([float("inf"), float("-inf"), float("nan"), 3e38, -3e38, 0.0, 1e-42, -1e-42], device="cuda")
    vmin[8], vmax[8] = 0.0, float("inf")            # divisor not finite: per-column IEEE fallback
    vmin[9], vmax[9] = -3e38, 3e38                  # range overflows
    rng = (vmax - vmin) + 1e-8
    ref = torch.nan_to_num(torch.round(((x - vmin) / rng) * lm1), nan=0.0, posinf=lm1, neginf=0.0).clamp(0, lm1)
    got = T.ops_sq.sq_encode(x, vmin, vmax, bits)
    if bits == 4:
        want = (ref[:, 0::2].to(torch.uint8) << 4) | ref[:, 1::2].to(torch.uint8)
        assert torch.equal(got, want)
    elif bits == 8:
        assert torch.equal(got, ref.to(torch.uint8))
    else:
        assert torch.equal(got.to(torch.int32) & 0xFFFF, ref.to(torch.int32))
    # decode: every level of every width against separate IEEE ops
    L = 1 << bits
    q = (torch.arange(L * 64, device="cuda") % L).reshape(-1, 64)
    vmn, vmx = vmin[:64].contiguous(), vmax[:64].contiguous()
    rr = (vmx - vmn) + 1e-8
    if bits == 4:
        codes = ((q[:, 0::2] << 4) | q[:, 1::2]).to(torch.uint8)
    elif bits == 8:
        codes = q.to(torch.uint8)
    else:
        codes = q.to(torch.int16)
    rec = T.ops_sq.sq_decode(codes, vmn, vmx, bits, 64)
    # tensor / tensor: torch turns tensor / python-scalar into a multiplication by the reciprocal
    want = (q.float() / torch.full_like(rr, lm1)) * rr + vmn
    assert torch.equal(torch.nan_to_num(rec, nan=-1.0), torch.nan_to_num(want, nan=-1.0))


def test_sq_minmax_accumulates_over_batches_and_saturates(T):
    X = make_data(3000, 64, 3, normalize=False)
    a = T.ops_sq.sq_minmax(T.dev(X[:1000]))
    a = T.ops_sq.sq_minmax(T.dev(X[1000:]), *a)
    assert np.array_equal(a[0].cpu().numpy(), X.min(0)) and np.array_equal(a[1].cpu().numpy(), X.max(0))
    out_of_range = T.dev(X * 10)
    codes = T.ops_sq.sq_encode(out_of_range, a[0], a[1], 8).cpu().numpy()
    assert codes.min() == 0 and codes.max() == 255  # documented divergence: saturate, never wrap
    empty = T.torch.empty((0, 64), dtype=T.torch.float32, device="cuda")
    assert T.ops_sq.sq_encode(empty, a[0], a[1], 8).shape == (0, 64)


@pytest.mark.parametrize("D,total,nseg", [(64, 256, 8), (3072, 3072 * 4, 8), (100, 333, 7), (48, 48 * 8, 3)])
def test_saq_bit_exact_vs_oracle(T, D, total, nseg):
    X = (make_data(1500, D, seed=D, normalize=False, aniso=False) * np.linspace(3, 0.05, D, dtype=np.float32))
    plan = osq.saq_plan(X, total_bits=total, allowed_bits=(0, 2, 4, 6, 8), n_segments=nseg)
    ref = osq.saq_compress(X[:300], plan)
    fill = np.where(plan["dim_bits"] == 0, plan["mean"], plan["vmin"]).astype(np.float32)
    bits, off = T.dev(plan["dim_bits"]), T.dev(plan["bit_offset"])
    codes = T.ops_sq.saq_encode(T.dev(X[:300]), T.dev(plan["vmin"]), T.dev(plan["vmax"]), bits, off,
                                plan["code_bytes"])
    assert np.array_equal(codes.cpu().numpy(), ref)
    rec = T.ops_sq.saq_decode(codes, T.dev(fill), T.dev(plan["vmax"]), bits, off, D).cpu().numpy()
    assert np.array_equal(rec, osq.saq_decompress(ref, plan))


def test_saq_odd_widths_and_uniform_mode(T):
    """Bit widths outside the CLI default set (1,3,5,7) exercise every byte-straddling case."""
    D = 40
    X = make_data(200, D, seed=1, normalize=False)
    for allowed, total in (((1, 3, 5, 7), D * 4), ((0, 8), D * 3), ((3,), D * 3)):
        plan = osq.saq_plan(X, total_bits=total, allowed_bits=allowed, n_segments=5)
        ref = osq.saq_compress(X, plan)
        bits, off = T.dev(plan["dim_bits"]), T.dev(plan["bit_offset"])
        codes = T.ops_sq.saq_encode(T.dev(X), T.dev(plan["vmin"]), T.dev(plan["vmax"]), bits, off,
                                    plan["code_bytes"])
        assert np.array_equal(codes.cpu().numpy(), ref), allowed
        fill = np.where(plan["dim_bits"] == 0, plan["mean"], plan["vmin"]).astype(np.float32)
        rec = T.ops_sq.saq_decode(codes, T.dev(fill), T.dev(plan["vmax"]), bits, off, D).cpu().numpy()
        assert np.array_equal(rec, osq.saq_decompress(ref, plan)), allowed


def test_saq_decode_trailing_zero_width_dimensions(T):
    """Dimensions of width 0 at the end of the row have bit_offset == total bits (one past the last code word)."""
    torch = T.torch
    D, n = 256, 3000
    g = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn((n, D), generator=g, device="cuda")
    vmin, vmax = T.ops_sq.sq_minmax(x)
    bits = np.repeat(np.array([8, 4, 0, 0], dtype=np.int32), D // 4)
    off = np.concatenate([[0], np.cumsum(bits)]).astype(np.int32)
    code_bytes = int(off[-1] + 7) // 8
    db, bo = T.dev(bits), T.dev(off)
    codes = T.ops_sq.saq_encode(x, vmin, vmax, db, bo, code_bytes)
    rec = T.ops_sq.saq_decode(codes, vmin, vmax, db, bo, D)
    assert torch.equal(rec[:, D // 2:], vmin[D // 2:].expand(n, -1))          # 0-bit dims decode to the fill value
    err = (rec[:, : D // 4] - x[:, : D // 4]).abs().max().item()
    assert err <= (vmax - vmin)[: D // 4].max().item() / 255 * 0.51
