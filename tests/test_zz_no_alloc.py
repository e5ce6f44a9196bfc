"""gsr_sort_pairs_static: the allocation-free variant of gsr_sort_pairs (scratch in a static device array inside
libgsr.so; no counterpart in the reference).  Called through the C ABI
and compared with the existing entry point: a sort is integer work, so the bar is bit-exact.  Two calls in a row share
the static scratch - the second one must not be disturbed by what the first left behind.  Collected last on purpose
(file name), passes alone in a fresh process and after the other tests.  `pytest -m gpu`."""
import ctypes as C

import numpy as np
import pytest

from gaussiansplattingviewer_b200 import _gsr

pytestmark = pytest.mark.gpu


def _keys(rng, n, kbits):
    if kbits == 32:
        k = rng.integers(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
        k[: n // 3] &= np.uint32(0xFF)  # many ties: stability shows
        return k
    k = rng.integers(0, 2 ** 63, size=n, dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=n, dtype=np.uint64)
    k[: n // 3] &= np.uint64(0xFFFF)
    return k


def _run_static(lib, torch, keys, vals, b0, b1):
    kb = keys.dtype.itemsize * 8
    tk = torch.from_numpy(keys.view(np.int32 if kb == 32 else np.int64).copy()).cuda()
    tv = torch.from_numpy(vals.view(np.int32).copy()).cuda()
    kt, vt = torch.empty_like(tk), torch.empty_like(tv)  # the caller's buffers; the library allocates nothing
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _gsr.check(lib.gsr_sort_pairs_static(kb, C.c_void_p(tk.data_ptr()), C.c_void_p(kt.data_ptr()), C.c_void_p(tv.data_ptr()),
                                         C.c_void_p(vt.data_ptr()), len(keys), b0, b1, stream))
    torch.cuda.synchronize()
    return tk.cpu().numpy().view(keys.dtype), tv.cpu().numpy().view(np.uint32)


def _run_handle(r, torch, keys, vals, b0, b1):
    kb = keys.dtype.itemsize * 8
    tk = torch.from_numpy(keys.view(np.int32 if kb == 32 else np.int64).copy()).cuda()
    tv = torch.from_numpy(vals.view(np.int32).copy()).cuda()
    r.sort_pairs(tk, tv, b0, b1)
    return tk.cpu().numpy().view(keys.dtype), tv.cpu().numpy().view(np.uint32)


def test_static_sort_matches_the_allocating_entry_point_on_two_calls_in_a_row():
    import torch
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    lib = _gsr.load()
    r = B200Renderer(64, 64)
    rng = np.random.default_rng(77)
    # back to back on the shared scratch: a large call, then smaller / differently shaped ones, then large again
    cases = [(32, 1_000_003, 0, 32), (32, 4097, 0, 32), (64, 700_001, 0, 64), (32, 300_000, 5, 19), (32, 1, 0, 32),
             (64, 3073, 8, 40), (32, 1_000_003, 0, 32)]
    for kbits, n, b0, b1 in cases:
        keys = _keys(rng, n, kbits)
        vals = rng.permutation(n).astype(np.uint32)
        k_s, v_s = _run_static(lib, torch, keys, vals, b0, b1)
        k_h, v_h = _run_handle(r, torch, keys, vals, b0, b1)
        assert np.array_equal(k_s, k_h), (kbits, n, b0, b1)
        assert np.array_equal(v_s, v_h), (kbits, n, b0, b1)
        # and both are THE stable sort on those bits
        digit = (keys >> keys.dtype.type(b0)) & keys.dtype.type((1 << (b1 - b0)) - 1) if b1 - b0 < kbits else keys
        want = np.argsort(digit, kind="stable")
        assert np.array_equal(v_s, vals[want]) and np.array_equal(k_s, keys[want]), (kbits, n, b0, b1)


def test_static_sort_limits_and_errors():
    import torch
    lib = _gsr.load()
    assert torch.cuda.is_available()
    torch.zeros(1, device="cuda")  # a current device / context
    n = _gsr.GSR_SORT_STATIC_MAX_PAIRS + 1
    one = torch.zeros(4, dtype=torch.int32, device="cuda")
    p = C.c_void_p(one.data_ptr())
    rc = lib.gsr_sort_pairs_static(32, p, p, p, p, n, 0, 32, None)  # rejected before anything is touched
    assert rc == -1 and b"GSR_SORT_STATIC_MAX_PAIRS" in lib.gsr_last_error()
    assert lib.gsr_sort_pairs_static(16, p, p, p, p, 4, 0, 16, None) == -1
    assert lib.gsr_sort_pairs_static(32, p, p, None, p, 4, 0, 32, None) == -1
    assert lib.gsr_sort_pairs_static(32, p, p, p, p, 0, 0, 32, None) == 0
    # the largest size that is accepted
    m = _gsr.GSR_SORT_STATIC_MAX_PAIRS
    rng = np.random.default_rng(78)
    keys = rng.integers(0, 2 ** 32, size=m, dtype=np.uint64).astype(np.uint32)
    vals = np.arange(m, dtype=np.uint32)
    k_s, v_s = _run_static(lib, torch, keys, vals, 0, 32)
    want = np.argsort(keys, kind="stable")
    assert np.array_equal(v_s, vals[want]) and np.array_equal(k_s, keys[want])
