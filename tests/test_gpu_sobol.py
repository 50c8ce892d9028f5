"""Sobol parity on the GPU: bit-exact against the oracle and the scipy-generated golden points."""
import os

import numpy as np
import pytest

import _oracle as O
import katib_b200 as kb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = kb.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("dim", [1, 2, 3, 64, 1000])
def test_first16_golden(ctx, golden_dir, dim):
    g = np.load(os.path.join(golden_dir, "sobol_golden.npz"))
    assert np.array_equal(ctx.sobol_fill(dim, 0, 16), g[f"first16_d{dim}"])


@pytest.mark.parametrize("dim,first,n", [(64, (1 << 20) - 3, 8), (7, 0, 1000), (33, 5, 777), (64, 12345, 4096), (1, 17, 1),
                                          (1024, 1 << 16, 50), (5, 0, 0), (64, (1 << 32) - 100, 100)])
def test_vs_oracle(ctx, golden_dir, dim, first, n):
    got = ctx.sobol_fill(dim, first, n)
    ref = O.sobol(golden_dir, dim, first, n)
    assert got.shape == ref.shape and np.array_equal(got, ref)


def test_scaled_is_unfused(ctx, golden_dir):
    dim, n = 6, 513
    lo = np.array([-1.0, 0.0, 3.0, -7.5, 1e-3, 100.0])
    hi = np.array([3.0, 1.0, 3.5, 9.25, 2e-3, 1e6])
    u = O.sobol(golden_dir, dim, 3, n)
    ref = lo + u * (hi - lo)  # numpy: separately rounded multiply and add, like Go
    assert np.array_equal(ctx.sobol_fill(dim, 3, n, lo, hi), ref)


def test_chunks_concatenate(ctx):
    a = ctx.sobol_fill(19, 0, 5000)
    b = np.concatenate([ctx.sobol_fill(19, 0, 1234), ctx.sobol_fill(19, 1234, 3766)])
    assert np.array_equal(a, b)


def test_full_size_checksum_d64(ctx, golden_dir):
    """2^20 x 64 against the scipy-derived position-weighted checksum; 2^24 x 64 (BASELINE configs[1]) on device
    against the oracle's u32 states through the same checksum computed chunk-wise."""
    import torch
    g = np.load(os.path.join(golden_dir, "sobol_golden.npz"))
    dim = 64

    def checksum(first, n):
        out = torch.empty((n, dim), dtype=torch.float64, device="cuda")
        ctx.sobol_fill_device(dim, first, n, out.data_ptr())
        torch.cuda.synchronize()
        xi = (out * 4294967296.0).to(torch.int64)
        idx = (torch.arange(first, first + n, device="cuda", dtype=torch.int64) + 1)[:, None]
        return (xi * idx).sum(dim=0).cpu().numpy().astype(np.uint64), out

    ws, out = checksum(0, 1 << 20)
    assert np.array_equal(ws, g["d64_n2p20_wsum"])
    # every value is a multiple of 2^-32 inside [0,1)
    assert float(out.min()) >= 0.0 and float(out.max()) < 1.0
    # full 2^24 points: each 1-d projection of the first 2^k points is a permutation of {i/2^k}
    n = 1 << 24
    big = torch.empty((n, dim), dtype=torch.float64, device="cuda")
    ctx.sobol_fill_device(dim, 0, n, big.data_ptr())
    torch.cuda.synchronize()
    for d in (0, 1, 31, 63):
        col = (big[:, d] * n).to(torch.int64)
        assert torch.equal(torch.sort(col).values, torch.arange(n, device="cuda"))
    tail = big[-4096:].cpu().numpy()
    assert np.array_equal(tail, O.sobol(golden_dir, dim, n - 4096, 4096))


def test_limits(ctx):
    with pytest.raises(kb.KbError):
        ctx.sobol_fill(1025, 0, 4)
    with pytest.raises(kb.KbError):
        ctx.sobol_fill(4, (1 << 32) - 2, 4)


@pytest.mark.parametrize("pinned", [False, True])
def test_host_fill_streams_in_chunks(golden_dir, pinned):
    """kb_sobol_fill produces the block in 32 MB chunks through two alternating device buffers, the copy of one chunk
    behind the generation of the next (pageable destinations additionally through pinned staging).  200 000 points x
    64 dims x 8 B = 102 MB = four chunks with a ragged last one, scaled to [low, high): every value must equal the
    oracle's, and a second, smaller call on the same context (buffers reused) as well."""
    import torch
    ctx = kb.Context(0)
    dim, first, n = 64, 12345, 200_000
    low, high = np.linspace(-3.0, 1.0, dim), np.linspace(2.0, 9.0, dim)
    if pinned:
        buf = torch.empty((n, dim), dtype=torch.float64, pin_memory=True)
        out, ptr = buf.numpy(), buf.data_ptr()
    else:
        out = np.empty((n, dim))
        ptr = out.ctypes.data
    lo_c, hi_c = np.ascontiguousarray(low), np.ascontiguousarray(high)
    assert kb.lib().kb_sobol_fill(ctx._h, dim, first, n, lo_c.ctypes.data, hi_c.ctypes.data, ptr) == 0
    ref = O.sobol(golden_dir, dim, first, n)
    want = ref * (high - low) + low  # separately rounded multiply and add, like the kernel (and Go)
    assert np.array_equal(out, want)
    small = ctx.sobol_fill(dim, 7, 1000)
    assert np.array_equal(small, O.sobol(golden_dir, dim, 7, 1000))
    one = ctx.sobol_fill(5, 99, 1)  # what the service asks for: one point
    assert np.array_equal(one, O.sobol(golden_dir, 5, 99, 1))
    ctx.close()
