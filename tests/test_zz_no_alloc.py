"""ecoz2b200_format_lines_noalloc: the device report formatter with its scratch in a fixed
__device__ array (no cudaMalloc / cudaFree in the call).  No counterpart in the reference; the
checker is the library's existing entry point ecoz2b200_debug_format_lines (which allocates) and,
through it, printf("%d,%f\\n") of report.c:118-121.

Collected last on purpose (file name).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _values(n, seed):
    rng = np.random.default_rng(seed)
    cw = rng.integers(0, 4096, size=n).astype(np.int32)
    md = np.abs(rng.standard_normal(n)) * 10.0 ** rng.integers(-9, 7, size=n)
    md[::7] = rng.random(len(md[::7]))                       # the usual range of min distortions
    md[::11] = np.round(md[::11], 6) + 5e-7                  # ties at the 7th digit
    md[::13] = -md[::13]                                     # (non gain-normalised frames)
    md[:4] = [0.0, -0.0, 5e-324, 2.0 ** 62 * (1 - 2.0 ** -52)]
    cw[:3] = [0, -1, 32767]
    return cw, md


def _run_noalloc(torch, vq, cw, md, capacity=None, stream=None):
    n = len(cw)
    d_cw = torch.from_numpy(cw).cuda()
    d_md = torch.from_numpy(md).cuda()
    cap = 40 * max(n, 1) if capacity is None else capacity
    d_text = torch.full((max(cap, 1),), 0x23, dtype=torch.uint8, device="cuda")
    d_bytes = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = vq.format_lines_noalloc(d_cw.data_ptr(), d_md.data_ptr(), n, d_text.data_ptr(), cap, d_bytes.data_ptr(),
                                 stream.cuda_stream if stream is not None else 0)
    if stream is not None:
        stream.synchronize()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    if rc != 0:
        return rc, None
    nb = int(d_bytes.item())
    if nb < 0:                                               # ~0: a value could not be formatted
        return rc, None
    return rc, bytes(d_text[:nb].cpu().numpy().tobytes())


def test_noalloc_formatter_equals_allocating_entry_point_twice(built_lib):
    import torch
    from ecoz2rs_b200 import vq
    torch.cuda.init()
    # two calls in a row share the scratch array: different sizes and contents, the second smaller
    # than the first so that stale block sums of the first call lie behind its own
    for n, seed in ((300_001, 1), (70_003, 2), (300_001, 1)):
        cw, md = _values(n, seed)
        rc_ref, want = vq.format_lines_on_device(cw, md)
        assert rc_ref == 0
        rc, got = _run_noalloc(torch, vq, cw, md)
        assert rc == 0, vq.last_error()
        assert got == want                                   # text: byte for byte
        assert got.count(b"\n") == n
    # spot check against Python's own %f (glibc rules)
    lines = got.split(b"\n")
    for v in (0, 1, 2, 3, 4, 1000, n - 1):
        assert lines[v] == (b"%d,%s" % (cw[v], ("%f" % md[v]).encode()))


def test_noalloc_formatter_on_a_side_stream_and_after_a_refused_value(built_lib):
    import torch
    from ecoz2rs_b200 import vq
    cw, md = _values(50_000, 3)
    bad = md.copy()
    bad[777] = np.inf
    rc, got = _run_noalloc(torch, vq, cw, bad)
    assert rc == 0 and got is None                           # *d_bytes = ~0: the caller falls back
    rc_ref, want = vq.format_lines_on_device(cw, md)
    s = torch.cuda.Stream()
    rc, got = _run_noalloc(torch, vq, cw, md, stream=s)      # the flag of the previous call is cleared
    assert rc_ref == 0 and rc == 0 and got == want


def test_noalloc_formatter_rejects_what_it_cannot_hold(built_lib):
    import torch
    from ecoz2rs_b200 import vq
    cw, md = _values(1000, 4)
    d = torch.zeros(16, dtype=torch.int64, device="cuda")
    # beyond the fixed scratch: refused before anything is touched
    rc = vq.format_lines_noalloc(d.data_ptr(), d.data_ptr(), vq.NOALLOC_MAX_LINES + 1, d.data_ptr(),
                                 40 * (vq.NOALLOC_MAX_LINES + 1), d.data_ptr())
    assert rc == vq.ERR_ARG and "exceed the maximum" in vq.last_error()
    # text buffer too small
    rc, _ = _run_noalloc(torch, vq, cw, md, capacity=40 * 1000 - 1)
    assert rc == vq.ERR_ARG and "text_capacity" in vq.last_error()
    # plain host memory is not a device buffer
    rc = vq.format_lines_noalloc(cw.ctypes.data, md.ctypes.data, 1000, d.data_ptr(), 40 * 1000, d.data_ptr())
    assert rc == vq.ERR_ARG and "device-accessible" in vq.last_error()
    # n = 0: *d_bytes = 0
    rc, got = _run_noalloc(torch, vq, cw[:0], md[:0])
    assert rc == 0 and got == b""
    # the maximum itself is accepted
    n = vq.NOALLOC_MAX_LINES
    cw, md = _values(n, 5)
    rc, got = _run_noalloc(torch, vq, cw, md)
    rc_ref, want = vq.format_lines_on_device(cw, md)
    assert rc == 0 and rc_ref == 0 and got == want
