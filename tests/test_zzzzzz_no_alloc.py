"""cusk_bed_marker_stats_noalloc (ci-gwas_b200/csrc/noalloc.cu): the allocation-free variant of the prep-bed statistics
against the existing entry point cusk_bed_marker_stats, through the C ABI, on two calls in a row (the second one shorter
than the first: the shared scratch keeps the first call's contents).  No counterpart in the reference.
Collected last on purpose.

Tolerance: the two entry points run the SAME two kernels on the same bytes, so means / stds must be bit-identical and the
integer outputs equal; the comparison below is exact.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


def _noalloc(lib, torch, bed, m, n, stream=None):
    d_bed = torch.from_numpy(np.ascontiguousarray(bed).reshape(-1)).cuda()
    d_means = torch.full((m,), -7.0, dtype=torch.float32, device="cuda")
    d_stds = torch.full((m,), -7.0, dtype=torch.float32, device="cuda")
    d_modes = torch.full((m,), -7, dtype=torch.int32, device="cuda")
    d_counts = torch.full((m, 4), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # torch's allocations and fills are done before the entry point is timed / checked
    before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    rc = lib.cusk_bed_marker_stats_noalloc(C.c_void_p(d_bed.data_ptr()), m, n, C.c_void_p(d_means.data_ptr()),
                                           C.c_void_p(d_stds.data_ptr()), C.c_void_p(d_modes.data_ptr()),
                                           C.c_void_p(d_counts.data_ptr()), C.c_void_p(stream or 0))
    assert rc == 0, lib.cusk_last_error().decode()
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == before
    return (d_means.cpu().numpy(), d_stds.cpu().numpy(), d_modes.cpu().numpy(), d_counts.cpu().numpy(),
            free_before - free_after)


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape
    if a.dtype.kind == "f":
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"max abs diff {np.nanmax(np.abs(a - b))}"
    else:
        assert np.array_equal(a.astype(np.int64), b.astype(np.int64))


def test_noalloc_equals_bed_marker_stats_two_calls_in_a_row(ctx):
    import torch
    from cigwas_b200 import synth

    lib = ctx.lib
    lib.cusk_last_error.restype = C.c_char_p
    cases = [(4099, 700, 11, 0.01), (1001, 37, 12, 0.0)]  # (samples, markers, stream, missing rate); second is SHORTER
    grew = []
    for n, m, stream, miss in cases:
        bed = synth.random_bed(n, m, stream=stream, missing_rate=miss)
        means, stds, modes, counts = ctx.bed_marker_stats(bed, m, n)  # the existing entry point (host buffers)
        got = _noalloc(lib, torch, bed, m, n)
        _same(got[0], means)
        _same(got[1], stds)
        _same(got[2], modes)
        _same(got[3], counts)
        assert (counts.sum(1) == n).all()
        grew.append(got[4])
    # the same input again on a stream of the caller's: same result (the scratch held the previous call's counts)
    n, m, stream, miss = cases[0]
    bed = synth.random_bed(n, m, stream=stream, missing_rate=miss)
    s = torch.cuda.Stream()
    again = _noalloc(lib, torch, bed, m, n, stream=s.cuda_stream)
    means, stds, modes, counts = ctx.bed_marker_stats(bed, m, n)
    _same(again[0], means)
    _same(again[1], stds)
    _same(again[3], counts)
    # no device memory was taken by the second call (the first may load the module's kernels lazily)
    assert grew[1] <= 0, grew


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(ctx):
    import torch

    lib = ctx.lib
    lib.cusk_last_error.restype = C.c_char_p
    limit = 65536  # CUSK_NOALLOC_MAX_MARKERS, include/cusk_b200.h
    x = torch.zeros(16, dtype=torch.uint8, device="cuda")
    p = C.c_void_p(x.data_ptr())
    rc = lib.cusk_bed_marker_stats_noalloc(p, limit + 1, 4, p, p, p, None, None)
    assert rc != 0
    msg = lib.cusk_last_error().decode()
    assert "CUSK_NOALLOC_MAX_MARKERS" in msg and str(limit + 1) in msg
    assert lib.cusk_bed_marker_stats_noalloc(None, 1, 4, p, p, p, None, None) != 0
    assert lib.cusk_bed_marker_stats_noalloc(p, 0, 4, p, p, p, None, None) != 0
