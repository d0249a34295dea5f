"""The allocation-free exclusive scan (sphb200_exclusive_scan_u32_static: scratch in a file-scope device array) against the
existing entry point (sphb200_exclusive_scan_u32), through the C ABI, on two calls in a row that share the scratch.
Integer work: bit-exact.  No counterpart in the reference.
This file is named to be collected last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 16384 * 4096  # the stated limit of the static scratch (include/sphb200.h)


def _scan_static(S, torch, values):
    n = len(values)
    d_in = torch.from_numpy(values.view(np.int32)).to("cuda:0")
    d_out = torch.full((n + 1,), -1, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()                       # the library works on its own stream
    S.exclusive_scan_static(d_in.data_ptr() if n else 0, d_out.data_ptr(), n)
    S.synchronize()
    out = d_out.cpu().numpy().view(np.uint32)
    return out[:n], int(out[n])


def test_static_scratch_scan_equals_the_allocating_scan_on_two_calls_in_a_row(pk, gpu):
    import torch
    S = gpu.SPHSystem(3, [0, 0, 0], [1, 1, 1], 0.1)
    rng = np.random.default_rng(2026)
    # the reference's known-answer vector first (test_exclusive_scan.cpp:8-25), through the new entry point
    kat = np.array([3, 2, 3, 5, 0, 1, 3, 2, 5, 1, 0], dtype=np.uint32)
    out, total = _scan_static(S, torch, kat)
    assert out.tolist() == [0, 3, 5, 8, 13, 13, 14, 17, 19, 24, 25] and total == 25
    # two calls in a row sharing the scratch: a large one (733 tiles, ragged tail), then a smaller one whose tiles are a
    # prefix of the first call's - stale tile sums behind them must not leak into the result
    for n in (3_000_001, 70_003):
        v = rng.integers(0, 1000, size=n, dtype=np.uint32)
        ref_out, ref_total = S.exclusive_scan(v)
        out, total = _scan_static(S, torch, v)
        assert total == ref_total == int(v.sum(dtype=np.uint64))
        assert np.array_equal(out, ref_out)
    # wrap-around arithmetic is the same too (u32 sums, like the reference's UnsignedInt scan in the float build)
    v = np.full(5000, 0xFFFFFFF0, dtype=np.uint32)
    ref_out, ref_total = S.exclusive_scan(v)
    out, total = _scan_static(S, torch, v)
    assert np.array_equal(out, ref_out) and total == ref_total
    # empty input: the total is written, nothing else
    out, total = _scan_static(S, torch, np.zeros(0, dtype=np.uint32))
    assert len(out) == 0 and total == 0
    S.close()


def test_static_scratch_scan_rejects_sizes_beyond_its_scratch(pk, gpu):
    import torch
    S = gpu.SPHSystem(3, [0, 0, 0], [1, 1, 1], 0.1)
    d = torch.zeros(16, dtype=torch.int32, device="cuda:0")
    with pytest.raises(Exception) as e:
        S.exclusive_scan_static(d.data_ptr(), d.data_ptr(), MAX_N + 1)   # rejected before anything is launched
    assert "exceeds the static scratch" in str(e.value) and str(MAX_N) in str(e.value)
    # the context is still usable afterwards
    out, total = _scan_static(S, torch, np.arange(10, dtype=np.uint32))
    assert out.tolist() == [0, 0, 1, 3, 6, 10, 15, 21, 28, 36] and total == 45
    S.close()
