"""sage_seq_mean_static: the allocation-free variant of sage_seq_mean (include/sage_b200.h; no counterpart in the
reference).  Its reduction scratch is a static device array inside
the library, shared by every call on a device and never cleared between calls except for what must start at zero.

Called here THROUGH THE C ABI (ctypes) and held to the existing entry point: same kernel, same arithmetic, so the
tolerance is the one the quantiser tests use for this path - BIT-EXACT (torch.equal).  Two calls in a row with
different shapes, so that the second call's arrival counters and partial sums land on the first call's leftovers.

This file is named to be collected last: a failure here must not hide the
other tests.  It passes alone in a fresh process and after the rest of the suite (it depends on no other test's
state: it makes its own set-up call, which is also what resolves the array's device address).
"""
import pytest
import torch

import sageattention_b200 as sb
from sageattention_b200 import _lib
from sageattention_b200.quant import _dims, _dtype_code, _stream

DEV = "cuda:0"
MAX_BYTES = 32 << 20  # SAGE_SEQ_MEAN_STATIC_MAX_BYTES


def _need(b, h, n, d):
    return b * h * ((n + 255) // 256) * 3 * d * 4 + b * h * 4 + 256


def _static_mean(x, layout="HND"):
    """x.mean(seq) through sage_seq_mean_static; the caller (this test) provides input and output only."""
    lib = _lib.load()
    b, h, n, d, x4 = _dims(x, layout)
    out = torch.empty((b, h, d), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.sage_seq_mean_static(x4, out.data_ptr(), b, h, n, d, _dtype_code(x), _stream(x))
    _lib.check(rc)
    return out


def _make(b, h, n, d, dtype, seed, layout="HND"):
    g = torch.Generator(device=DEV).manual_seed(seed)
    x = (torch.randn(b, h, n, d, generator=g, device=DEV) + 1.5 * torch.randn(1, h, 1, d, generator=g, device=DEV)).to(dtype)
    return x.transpose(1, 2).contiguous() if layout == "NHD" else x


@pytest.mark.gpu
def test_static_scratch_matches_workspace_entry_point_on_two_calls_in_a_row():
    # set-up call (resolves the array's device address once for this device)
    _static_mean(_make(1, 1, 64, 64, torch.float16, 0))
    torch.cuda.synchronize()
    # first call: many chunks and (b, h) slices -> fills a large part of the array with partial sums
    x1 = _make(4, 32, 4096, 128, torch.bfloat16, 1)            # BASELINE C2's K: 3.1 MB of scratch
    # second call: a different decomposition, so its counters sit where call 1 left partial sums, and its
    # partials where call 1 left other partials; NHD layout and fp16 for good measure
    x2 = _make(2, 6, 1000, 64, torch.float16, 2, layout="NHD")
    ref1 = sb.seq_mean(x1)                                       # existing entry point (workspace from the caller)
    ref2 = sb.seq_mean(x2, tensor_layout="NHD")
    out_keep = torch.empty_like(ref1)                            # (allocate everything this test needs up front)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0 = torch.cuda.memory_allocated()
    lib = _lib.load()
    b, h, n, d, x4 = _dims(x1, "HND")
    _lib.check(lib.sage_seq_mean_static(x4, out_keep.data_ptr(), b, h, n, d, _dtype_code(x1), _stream(x1)))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    # no allocation on either side of the ABI: the driver's free memory and torch's allocator are where they were
    assert free1 == free0 and torch.cuda.memory_allocated() == alloc0
    assert torch.equal(out_keep, ref1)
    got2 = _static_mean(x2, "NHD")
    torch.cuda.synchronize()
    assert torch.equal(got2, ref2)
    # and once more with the first input: the array now holds call 2's leftovers
    assert torch.equal(_static_mean(x1), ref1)
    # the largest named shape that must fit (Wan2.1 K: B1 H40 N75600 D128, 18.2 MB of the 32 MiB)
    assert _need(1, 40, 75600, 128) <= MAX_BYTES
    x3 = _make(1, 40, 75600, 128, torch.bfloat16, 3)
    assert torch.equal(_static_mean(x3), sb.seq_mean(x3))


def test_static_scratch_rejects_sizes_beyond_the_array():
    """The limit is checked before anything is launched or resolved (so this needs no GPU): -1 and a message
    naming both sizes."""
    lib = _lib.load()
    b, h, n, d = 8, 64, 131072, 128
    assert _need(b, h, n, d) > MAX_BYTES
    dummy = torch.zeros(16, dtype=torch.float16)   # never dereferenced: the call fails before any launch
    x4 = _lib.Tensor4(dummy.data_ptr(), h * n * d, n * d, d)
    rc = lib.sage_seq_mean_static(x4, dummy.data_ptr(), b, h, n, d, _lib.DTYPE_F16, None)
    assert rc == -1
    msg = lib.sage_last_error().decode()
    assert "sage_seq_mean_static" in msg and str(_need(b, h, n, d)) in msg and str(MAX_BYTES) in msg
    with pytest.raises(ValueError):
        _lib.check(rc)
