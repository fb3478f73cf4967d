"""GPU: ms_kde_select_static -- the allocation-free variant of ms_kde_select (scratch in a static device array inside
the library; no counterpart in the reference) -- against ms_kde_select
through the C ABI, two calls in a row.  Both entry points launch the same two kernels, so the results are equal to the
last bit; the assertion uses the densities' bar of the other parity tests (1e-9 relative / 1e-12 absolute).

Collected last on purpose: a failure here hides nothing else."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _select(L, _lib, raw, nv, nlabels, select, nsel, get_counts, static, dev):
    import torch
    o64 = torch.full((nv,), float('nan'), dtype=torch.float64, device=dev)
    o32 = torch.full((nv,), float('nan'), dtype=torch.float32, device=dev)
    sel = np.asarray(select, np.uint8).tobytes()
    st = _lib.stream_ptr(dev)
    if static:
        rc = L.ms_kde_select_static(raw.data_ptr(), nv, nlabels, sel, float(nsel), int(get_counts),
                                    o64.data_ptr(), o32.data_ptr(), st)
        _lib.check(rc, 'ms_kde_select_static')
    else:
        ws = _lib.workspace(L.ms_kde_select_workspace_bytes(nv), dev)
        rc = L.ms_kde_select(raw.data_ptr(), nv, nlabels, sel, float(nsel), int(get_counts),
                             o64.data_ptr(), o32.data_ptr(), ws.data_ptr(), ws.numel(), st)
        _lib.check(rc, 'ms_kde_select')
    torch.cuda.synchronize(dev)
    return o64.cpu().numpy(), o32.cpu().numpy()


def test_kde_select_static_matches_kde_select_twice():
    import torch
    from memsurfer_b200 import _lib
    _lib.require_cuda()
    L = _lib.lib()
    dev = torch.device('cuda:0')
    gen = torch.Generator(device='cpu').manual_seed(7)
    nlabels = 3
    # two calls in a row on the shared scratch: the larger problem first, so that the second call runs over the
    # first one's left-overs (partial sums of 196 blocks, densities beyond nv)
    cases = [(50000, [1, 0, 1], 31234.0, 1), (12345, [0, 1, 0], 4000.0, 1), (50000, [1, 1, 1], 50000.0, 0)]
    for nv, select, nsel, get_counts in cases:
        raw = torch.rand((nlabels, nv), dtype=torch.float64, generator=gen).to(dev)
        want64, want32 = _select(L, _lib, raw, nv, nlabels, select, nsel, get_counts, False, dev)
        got64, got32 = _select(L, _lib, raw, nv, nlabels, select, nsel, get_counts, True, dev)
        assert np.isfinite(got64).all() and np.isfinite(got32).all()
        assert (np.abs(got64 - want64) <= 1e-9 * np.abs(want64) + 1e-12).all(), f"nv {nv}: fp64 densities differ"
        assert np.array_equal(got32, want32), f"nv {nv}: fp32 densities differ"
        if get_counts:
            assert abs(got64.sum() - nsel) <= 1e-9 * nsel


def test_kde_select_static_rejects_sizes_beyond_its_scratch():
    import torch
    from memsurfer_b200 import _lib
    _lib.require_cuda()
    L = _lib.lib()
    dev = torch.device('cuda:0')
    raw = torch.zeros(16, dtype=torch.float64, device=dev)
    out = torch.zeros(16, dtype=torch.float64, device=dev)
    too_many = 1048576 + 1                                       # MS_KDE_SELECT_STATIC_MAX_NV + 1 (include/memsurfer_b200.h)
    rc = L.ms_kde_select_static(raw.data_ptr(), too_many, 1, bytes([1]), 1.0, 0, out.data_ptr(), None, _lib.stream_ptr(dev))
    assert rc != 0
    msg = L.ms_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert 'MS_KDE_SELECT_STATIC_MAX_NV' in msg and 'ms_kde_select_static' in msg
    torch.cuda.synchronize(dev)
    assert float(out.abs().sum()) == 0.0                         # nothing was launched
