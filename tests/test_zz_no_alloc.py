"""gsm_observed_hist_static: the observed-histogram kernel for hosts that must not allocate.

Some processes that embed libgsmfret.so cannot let it call cudaMalloc: their device memory is
already committed to a pool of their own, or they sit in a loop that cannot absorb the device-wide
synchronisation an allocation or a free brings, or they are workers that only ever receive buffers
and a stream.  For them the library offers one entry point whose working set (the 64-bit bin
counters) is an array compiled into the library instead of a cudaMalloc'ed block.  This file checks,
through the C ABI, that it returns exactly what the allocating path (gsm_init_data +
gsm_observed_hist) returns, call after call although all calls share that one array, that it takes
no device memory, and that it refuses an axis larger than the array.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here
must not hide the parity tests.  The tests need no state from other files and leave none behind.
"""
import numpy as np
import pytest

from gsmfretda_b200 import synth

pytestmark = pytest.mark.gpu

GSM_STATIC_MAX_BINS = 4096
GSM_EINVAL, GSM_ERANGE = -1, -6


def _allocating_path(gsm, b, n_bins):
    """What the existing entry points report for these bursts: the comparison value."""
    with gsm.Mc(0, 1) as mc:
        mc.set_hist(n_bins, 0)          # the left pad only matters to chi-square
        mc.init_bursts(b)
        return mc.observed_hist()


def _static_path(lib, torch, sg_dev, n_bins, out_dev, stream=None):
    out_dev.fill_(-1.0)                       # stale output must not pass for a result
    s = stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream
    rc = lib.gsm_observed_hist_static(sg_dev.data_ptr(), sg_dev.numel(), n_bins, out_dev.data_ptr(), s)
    assert rc == 0, lib.gsm_last_error()
    torch.cuda.synchronize()
    return out_dev.cpu().numpy()


def _bits(x):
    return np.ascontiguousarray(x, np.float32).view(np.uint32)


def test_static_scratch_equals_allocating_path_on_consecutive_calls(gsm):
    """Two calls in a row on different data (the second would inherit the first one's counts if the
    shared array were not cleared), then the first data again; each bit-equal to gsm_observed_hist,
    the tolerance every histogram test of this suite uses (integer counts, exactly representable)."""
    import torch
    torch.cuda.set_device(0)
    lib = gsm.load()
    n_bins = 200
    b1 = synth.workload(2, n_bursts=5000, R=1).bursts()
    b2 = synth.workload(5, n_bursts=1777, R=1).bursts()
    # outside [0, 1) and NaN go to the outer bins and must not leak into the inner ones
    b2.SgDivSr[:7] = np.array([1.0, 1.5, -0.25, np.nan, 0.0, np.nextafter(np.float32(1), np.float32(0)), -0.0], np.float32)
    want1 = _allocating_path(gsm, b1, n_bins)
    want2 = _allocating_path(gsm, b2, n_bins)
    assert want1.sum() > 0 and want2.sum() > 0 and not np.array_equal(want1, want2[: want1.size])

    sg1 = torch.from_numpy(np.ascontiguousarray(b1.SgDivSr, np.float32)).cuda()
    sg2 = torch.from_numpy(np.ascontiguousarray(b2.SgDivSr, np.float32)).cuda()
    out = torch.empty(n_bins, dtype=torch.float32, device="cuda")
    side = torch.cuda.Stream()
    torch.cuda.synchronize()

    got1 = _static_path(lib, torch, sg1, n_bins, out, stream=side)
    assert np.array_equal(_bits(got1), _bits(want1))
    free_before = torch.cuda.mem_get_info()[0]
    got2 = _static_path(lib, torch, sg2, n_bins, out, stream=side)
    assert np.array_equal(_bits(got2), _bits(want2))
    got1_again = _static_path(lib, torch, sg1, n_bins, out)      # and on the stream torch itself works on
    assert np.array_equal(_bits(got1_again), _bits(want1))
    # once the first call has resolved the array and loaded the kernels, a call takes nothing from the device
    assert torch.cuda.mem_get_info()[0] >= free_before


def test_static_scratch_other_axes_and_empty_input(gsm):
    import torch
    torch.cuda.set_device(0)
    lib = gsm.load()
    b = synth.workload(3, n_bursts=3001, R=1).bursts()
    sg = torch.from_numpy(np.ascontiguousarray(b.SgDivSr, np.float32)).cuda()
    for n_bins in (1000, 1, 37):              # wide first: a narrower axis afterwards sees cleared counters only
        out = torch.empty(n_bins, dtype=torch.float32, device="cuda")
        got = _static_path(lib, torch, sg, n_bins, out)
        assert np.array_equal(_bits(got), _bits(_allocating_path(gsm, b, n_bins))), n_bins
    # the largest axis the built-in array holds (beyond gsm_set_hist's 1022): every burst inside [0, 1) is counted once
    out = torch.empty(GSM_STATIC_MAX_BINS, dtype=torch.float32, device="cuda")
    got = _static_path(lib, torch, sg, GSM_STATIC_MAX_BINS, out)
    x = b.SgDivSr.astype(np.float64)
    inside = (x >= 0.0) & (x < 1.0)
    want = np.bincount((x[inside] * GSM_STATIC_MAX_BINS).astype(np.int64), minlength=GSM_STATIC_MAX_BINS)
    assert np.array_equal(got, want.astype(np.float32))
    # no bursts: zeros, whatever the previous call left
    out = torch.empty(200, dtype=torch.float32, device="cuda")
    out.fill_(-1.0)
    assert lib.gsm_observed_hist_static(None, 0, 200, out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert not out.cpu().numpy().any()


def test_static_scratch_refuses_what_it_cannot_hold(gsm):
    import torch
    torch.cuda.set_device(0)
    lib = gsm.load()
    sg = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = torch.full((GSM_STATIC_MAX_BINS + 1,), -1.0, dtype=torch.float32, device="cuda")
    rc = lib.gsm_observed_hist_static(sg.data_ptr(), 16, GSM_STATIC_MAX_BINS + 1, out.data_ptr(), None)
    assert rc == GSM_ERANGE
    msg = lib.gsm_last_error().decode()
    assert "GSM_STATIC_MAX_BINS" in msg and str(GSM_STATIC_MAX_BINS) in msg
    assert lib.gsm_observed_hist_static(sg.data_ptr(), 16, 0, out.data_ptr(), None) == GSM_EINVAL
    assert lib.gsm_observed_hist_static(sg.data_ptr(), -1, 200, out.data_ptr(), None) == GSM_EINVAL
    assert lib.gsm_observed_hist_static(sg.data_ptr(), 16, 200, None, None) == GSM_EINVAL
    assert lib.gsm_observed_hist_static(None, 16, 200, out.data_ptr(), None) == GSM_EINVAL
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == -1.0).all()   # a refused call launched nothing
