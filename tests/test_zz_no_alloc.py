"""ss_sort_pairs_static: the sort for callers that may not allocate device memory (a render thread on a
fixed memory budget, a plug-in whose host application owns the allocator).  Its scratch is an array
built into the library, so it must give exactly what ss_sort_pairs gives with caller-provided scratch.
Checked twice in a row because the built-in array still holds the first call's state when the second
call starts.  Collected last on purpose (file name), and self-contained so it also passes alone."""
import ctypes as C

import numpy as np
import pytest

from vulkanengine_b200 import scenesubmit as S

pytestmark = pytest.mark.gpu

STATIC_MAX_N = 1 << 20


def _dev(t):
    return C.c_void_p(t.data_ptr())


def _sort_both(lib, torch, keys_np, begin_bit, end_bit):
    """Runs the existing entry point and the allocation-free one on the same input; returns both results."""
    n = len(keys_np)
    dev = torch.device("cuda", 0)
    results = []
    for which in ("scratch", "static"):
        ka = torch.from_numpy(keys_np.view(np.int64).copy()).to(dev)
        va = torch.arange(n, dtype=torch.int32, device=dev)
        kb = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
        vb = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        if n == 0:
            ka = torch.empty(1, dtype=torch.int64, device=dev)
            va = torch.empty(1, dtype=torch.int32, device=dev)
        in_b = C.c_int(-1)
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if which == "scratch":
            sb = lib.ss_sort_scratch_bytes(C.c_uint64(n))
            scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
            st = lib.ss_sort_pairs(stream, _dev(ka), _dev(va), _dev(kb), _dev(vb), C.c_uint64(n), C.c_int(begin_bit),
                                   C.c_int(end_bit), _dev(scratch), C.c_size_t(sb), C.byref(in_b))
        else:
            st = lib.ss_sort_pairs_static(stream, _dev(ka), _dev(va), _dev(kb), _dev(vb), C.c_uint64(n),
                                          C.c_int(begin_bit), C.c_int(end_bit), C.byref(in_b))
        assert st == S.SS_OK, lib.ss_last_error(None).decode()
        torch.cuda.synchronize()
        k, v = (kb, vb) if in_b.value else (ka, va)
        results.append((k[:n].cpu().numpy().view(np.uint64), v[:n].cpu().numpy().view(np.uint32)))
    return results


@pytest.mark.parametrize("n,begin_bit,end_bit", [(200_003, 0, 64), (STATIC_MAX_N, 0, 19), (4097, 8, 29)])
def test_static_scratch_sort_matches_sort_with_caller_scratch(n, begin_bit, end_bit):
    import torch
    lib = S.load_library()
    rng = np.random.default_rng(n + end_bit)
    for call in range(2):  # two calls in a row: the second starts from the first one's leftovers
        keys = rng.integers(0, 2**64, size=n, dtype=np.uint64)
        (k_ref, v_ref), (k_new, v_new) = _sort_both(lib, torch, keys, begin_bit, end_bit)
        assert np.array_equal(k_new, k_ref), f"call {call}: keys differ"
        assert np.array_equal(v_new, v_ref), f"call {call}: values differ"
        sub = (keys >> np.uint64(begin_bit)) & np.uint64((1 << (end_bit - begin_bit)) - 1 if end_bit - begin_bit < 64
                                                          else 0xFFFFFFFFFFFFFFFF)
        order = np.argsort(sub, kind="stable")
        assert np.array_equal(v_new, order.astype(np.uint32)), f"call {call}: not the stable order"


def test_static_scratch_sort_rejects_sizes_beyond_its_maximum():
    import torch
    lib = S.load_library()
    dev = torch.device("cuda", 0)
    n = STATIC_MAX_N + 1
    ka = torch.zeros(n, dtype=torch.int64, device=dev)
    va = torch.zeros(n, dtype=torch.int32, device=dev)
    kb, vb = torch.empty_like(ka), torch.empty_like(va)
    in_b = C.c_int(-1)
    st = lib.ss_sort_pairs_static(None, _dev(ka), _dev(va), _dev(kb), _dev(vb), C.c_uint64(n), C.c_int(0),
                                  C.c_int(64), C.byref(in_b))
    assert st == S.SS_ERR_CAPACITY
    assert b"SS_SORT_STATIC_MAX_N" in lib.ss_last_error(None)
    st = lib.ss_sort_pairs_static(None, None, _dev(va), _dev(kb), _dev(vb), C.c_uint64(8), C.c_int(0), C.c_int(64),
                                  C.byref(in_b))
    assert st == S.SS_ERR_INVALID


def test_static_scratch_sort_on_a_second_device_and_stream_ownership():
    """The built-in scratch exists once per GPU that loads the library and its address is looked up per device: a
    sort on device 1 must work after sorts on device 0, and a stream of another device than the current one is
    refused instead of running kernels where the looked-up address means nothing.  Needs two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    lib = S.load_library()
    n = 50_000
    keys = np.random.default_rng(3).integers(0, 2**64, size=n, dtype=np.uint64)
    with torch.cuda.device(1):
        dev = torch.device("cuda", 1)
        ka = torch.from_numpy(keys.view(np.int64).copy()).to(dev)
        va = torch.arange(n, dtype=torch.int32, device=dev)
        kb, vb = torch.empty_like(ka), torch.empty_like(va)
        in_b = C.c_int(-1)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        st = lib.ss_sort_pairs_static(stream, _dev(ka), _dev(va), _dev(kb), _dev(vb), C.c_uint64(n), C.c_int(0), C.c_int(64),
                                      C.byref(in_b))
        assert st == S.SS_OK, lib.ss_last_error(None).decode()
        v = (vb if in_b.value else va).cpu().numpy().view(np.uint32)
        assert np.array_equal(v, np.argsort(keys, kind="stable").astype(np.uint32))
        other = torch.cuda.Stream(device=0)
        st = lib.ss_sort_pairs_static(C.c_void_p(other.cuda_stream), _dev(ka), _dev(va), _dev(kb), _dev(vb), C.c_uint64(n),
                                      C.c_int(0), C.c_int(64), C.byref(in_b))
        assert st == S.SS_ERR_INVALID
