"""k8sgpu_exclusive_scan_noalloc: the prefix sum for callers that may not allocate device
memory while running (a controller confined to a fixed memory slice next to the jobs it
schedules; code that cannot take the device-wide stall of cudaMalloc/cudaFree).

It keeps its block sums in a fixed array inside the library instead of a cudaMalloc'd buffer.
The test drives it through the C-ABI with caller-owned device buffers and compares it, bit for
bit, with the allocating entry point k8sgpu_debug_exclusive_scan -- twice in a row, the second
time on a shorter input, so that the leftovers of the first call sit in the shared array when
the second one runs.  pytest collects this file last (test_zz_*), so a failure here cannot
hide any other test from a `-x` run; it passes on its own and after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def capi():
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi as c

    return c


def run_noalloc(ctx, torch, a):
    """One call through the C-ABI on the current torch stream; returns (out, total) as numpy/int."""
    d_in = torch.from_numpy(a.astype(np.int32)).cuda()  # same bits as uint32
    d_out = torch.empty_like(d_in)
    d_tot = torch.zeros(1, dtype=torch.int32, device="cuda")
    ctx.exclusive_scan_noalloc(d_in.data_ptr(), a.size, d_out.data_ptr(), d_tot.data_ptr(),
                               torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return d_out.cpu().numpy().view(np.uint32), int(d_tot.cpu().numpy().view(np.uint32)[0])


def test_two_calls_in_a_row_match_the_allocating_entry_point(capi):
    import torch

    rng = np.random.default_rng(2024)
    with capi.Context(0) as ctx:
        # first call: 1.2 M elements -> 586 block sums written into the shared scratch
        # second call: 70 k elements -> only 35 of them are rewritten; the rest are stale
        for n in (1_200_003, 70_001):
            a = rng.integers(0, 5000, n, dtype=np.uint32)
            got, got_total = run_noalloc(ctx, torch, a)
            want, want_total = ctx.exclusive_scan(a)  # the existing, allocating entry point
            assert np.array_equal(got, want)          # integer work: bit exact
            assert got_total == want_total == int(a.sum(dtype=np.uint64) & 0xFFFFFFFF)
            ref = np.cumsum(a, dtype=np.uint64)       # and both against numpy itself
            assert got[0] == 0 and np.array_equal(got[1:], (ref[:-1] & 0xFFFFFFFF).astype(np.uint32))


def test_largest_size_against_numpy_and_no_allocation(capi):
    """n = 2^24 exactly: the one size that fills the built-in scratch to its last word (8192 block
    sums), where an off-by-one would run past the array.  Checked against numpy.cumsum directly,
    and the device's free memory is the same before and after the calls: nothing was allocated."""
    import torch

    n = capi.SCAN_NOALLOC_MAX_N
    rng = np.random.default_rng(7)
    with capi.Context(0) as ctx:
        a = rng.integers(0, 200, n, dtype=np.uint32)
        d_in = torch.from_numpy(a.view(np.int32)).cuda()
        d_out = torch.empty_like(d_in)
        d_tot = torch.zeros(1, dtype=torch.int32, device="cuda")
        stream = torch.cuda.current_stream().cuda_stream
        ctx.exclusive_scan_noalloc(d_in.data_ptr(), 16, d_out.data_ptr(), d_tot.data_ptr(), stream)  # set-up call
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        for m in (n, n - 1, n - 2047):  # full, one short of full, last block nearly empty
            ctx.exclusive_scan_noalloc(d_in.data_ptr(), m, d_out.data_ptr(), d_tot.data_ptr(), stream)
            torch.cuda.synchronize()
            want = np.cumsum(a[:m], dtype=np.uint64)
            got = d_out[:m].cpu().numpy().view(np.uint32)
            assert got[0] == 0 and np.array_equal(got[1:], (want[:-1] & 0xFFFFFFFF).astype(np.uint32))
            assert int(d_tot.cpu().numpy().view(np.uint32)[0]) == int(want[-1] & 0xFFFFFFFF)
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before


def test_small_inputs_in_place_and_empty(capi):
    import torch

    with capi.Context(0) as ctx:
        for n in (0, 1, 2048, 2049):
            a = np.arange(1, n + 1, dtype=np.uint32)
            got, tot = run_noalloc(ctx, torch, a)
            want, want_tot = ctx.exclusive_scan(a)
            assert np.array_equal(got, want) and tot == want_tot
        # d_out == d_in is allowed
        a = np.full(10_000, 3, dtype=np.uint32)
        buf = torch.from_numpy(a.astype(np.int32)).cuda()
        ctx.exclusive_scan_noalloc(buf.data_ptr(), a.size, buf.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(buf.cpu().numpy().view(np.uint32), ctx.exclusive_scan(a)[0])


def test_sizes_past_the_built_in_scratch_are_refused(capi):
    import torch

    with capi.Context(0) as ctx:
        buf = torch.zeros(16, dtype=torch.int32, device="cuda")
        with pytest.raises(capi.K8sGpuError) as e:  # rejected before anything is launched or read
            ctx.exclusive_scan_noalloc(buf.data_ptr(), capi.SCAN_NOALLOC_MAX_N + 1, buf.data_ptr())
        assert e.value.code == capi.E_CAPACITY
        assert str(capi.SCAN_NOALLOC_MAX_N) in str(e.value)
