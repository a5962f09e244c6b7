"""b200ad_video_pair_sums_static -- the no-allocation variant of b200ad_video_pair_sums (include/b200ad.h; scratch in a
file-scope device array of csrc/metrics.cu, address taken with cudaGetSymbolAddress) -- against the existing entry
point, through the C-ABI, on two calls in a row.  The sums are exact integers held in doubles, so the comparison is
bit-exact (the tolerance tests/test_metrics_gpu.py uses for them).

Collected last on purpose (file name).  -m gpu."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _frames(F, H, W, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (F, H, W, 3), dtype=torch.uint8, generator=g)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _workspace_entry(L, fr):
    """The existing entry point: caller-provided workspace."""
    from b200ad import lib as _lib
    F, n = fr.shape[0], fr[0].numel()
    ws = torch.empty(L.b200ad_metric_blocks() * 3 * (F - 1), device="cuda", dtype=torch.float64)
    out = torch.full((F - 1, 2), -1.0, device="cuda", dtype=torch.float64)
    _lib.check(L.b200ad_video_pair_sums(fr.data_ptr(), F, n, ws.data_ptr(), out.data_ptr(), _stream()),
               "b200ad_video_pair_sums")
    torch.cuda.synchronize()
    return out.cpu()


def _static_entry(L, fr, out):
    """The new entry point: inputs and the output only; `out` is allocated by the caller beforehand."""
    from b200ad import lib as _lib
    F, n = fr.shape[0], fr[0].numel()
    _lib.check(L.b200ad_video_pair_sums_static(fr.data_ptr(), F, n, out.data_ptr(), _stream()),
               "b200ad_video_pair_sums_static")
    torch.cuda.synchronize()
    return out[: 2 * (F - 1)].view(F - 1, 2).cpu()


def _numpy_sums(fr):
    x = fr.numpy().astype(np.int64).reshape(fr.shape[0], -1)
    sq = ((x[:-1] - x[1:]) ** 2).sum(1)
    fl = np.zeros_like(sq)
    fl[:-1] = np.abs(x[:-2] - 2 * x[1:-1] + x[2:]).sum(1)
    return np.stack([sq, fl], 1).astype(np.float64)


def test_static_scratch_entry_matches_the_workspace_entry_twice_in_a_row():
    from b200ad import lib as _lib
    L = _lib.load()
    assert L.b200ad_static_max_frames() == 64
    # every device buffer is created BEFORE the calls under test: the entry point itself allocates nothing
    cases = [_frames(16, 64, 96, seed=1).cuda(), _frames(5, 40, 72, seed=2).cuda(), _frames(64, 32, 32, seed=3).cuda()]
    out = torch.empty(2 * 63, device="cuda", dtype=torch.float64)
    for fr in cases[:2]:                               # first use of everything the timed loop touches: CUDA loads
        out.fill_(-1.0)                                # modules lazily (torch's fill / copy kernels and this library's),
        _static_entry(L, fr, out)                      # and the entry point looks its scratch address up once per device
    free_before = torch.cuda.mem_get_info()[0]
    got = []
    for fr in cases:                                   # back to back; the 2nd has fewer frames than the 1st, so the
        out.fill_(-1.0)                                # scratch still holds stale partials of the 1st beyond its rows
        got.append(_static_entry(L, fr, out).clone())
    assert torch.cuda.mem_get_info()[0] >= free_before, "the static entry point must not take device memory"
    for fr, g in zip(cases, got):
        ref = _workspace_entry(L, fr)
        assert torch.equal(g, ref), "static-scratch result differs from b200ad_video_pair_sums"
        assert np.array_equal(g.numpy(), _numpy_sums(fr.cpu())), "pair sums differ from exact integer arithmetic"
    # and the same input again gives the same answer (scratch contents of earlier calls do not leak in)
    out.fill_(-1.0)
    assert torch.equal(_static_entry(L, cases[0], out), got[0])


def test_static_scratch_entry_rejects_sizes_beyond_its_limit():
    from b200ad import lib as _lib
    L = _lib.load()
    fr = _frames(65, 16, 16, seed=4).cuda()
    out = torch.full((2 * 64,), -7.0, device="cuda", dtype=torch.float64)
    rc = L.b200ad_video_pair_sums_static(fr.data_ptr(), 65, fr[0].numel(), out.data_ptr(), _stream())
    assert rc == -1
    msg = L.b200ad_last_error().decode()
    assert "64" in msg and "limit" in msg, msg
    with pytest.raises(_lib.B200ADError):
        _lib.check(rc, "b200ad_video_pair_sums_static")
    torch.cuda.synchronize()
    assert bool((out == -7.0).all()), "a rejected call must not touch the output"
    # the workspace entry point still takes that size
    ref = _workspace_entry(L, fr)
    assert np.array_equal(ref.numpy(), _numpy_sums(fr.cpu()))


def test_static_scratch_entry_runs_under_stream_capture():
    """cudaMalloc is illegal while a stream is being captured (torch captures in the "global" error mode, where an
    allocation from this thread fails the capture): the entry point records into a CUDA graph after its one-time
    set-up call, and the replayed graph gives the workspace entry point's sums."""
    from b200ad import lib as _lib
    L = _lib.load()
    fr = _frames(12, 48, 80, seed=5).cuda()
    out = torch.full((2 * 63,), -1.0, device="cuda", dtype=torch.float64)
    ref = _workspace_entry(L, fr)
    _static_entry(L, fr, out)                          # set-up call outside the capture (address look-up, module load)
    out.fill_(-1.0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    F, n = fr.shape[0], fr[0].numel()
    with torch.cuda.graph(g):
        rc = L.b200ad_video_pair_sums_static(fr.data_ptr(), F, n, out.data_ptr(), _stream())
    assert rc == 0, L.b200ad_last_error()
    torch.cuda.synchronize()
    assert bool((out == -1.0).all()), "capture must record, not run"
    for _ in range(2):
        out.fill_(-1.0)
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out[: 2 * (F - 1)].view(F - 1, 2).cpu(), ref)
