"""b200_sort_pairs_k32_static — the allocation-free variant of the depth-key sort (scratch in a file-scope device array
inside the library; no counterpart in the reference) — against the
existing entry point b200_sort_pairs_k32, through the C ABI.  Integer work: the bar is bit-exact, like every other sort
test here.  Collected last on purpose (file name), runs alone or after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(n, end_bit, seed):
    g = torch.Generator().manual_seed(seed)
    keys = torch.randint(0, 2 ** end_bit, (n,), generator=g, dtype=torch.int64)
    keys[: n // 3] &= 0x3F  # long runs of equal keys: stability matters
    wrap = keys.clone()
    wrap[wrap >= (1 << 31)] -= (1 << 32)  # two's-complement storage of the unsigned key
    return keys, wrap.to(torch.int32), torch.arange(n, dtype=torch.int32)


def _run(lib, L, static, k32, vals, n, end_bit):
    ka, va = k32.to(DEV), vals.to(DEV)
    kb, vb = torch.empty_like(ka), torch.empty_like(va)
    in_b = C.c_int(-1)
    if static:
        # everything the call needs is allocated above, by the caller; the entry point itself allocates nothing
        L.check(lib.b200_sort_pairs_k32_static(ka.data_ptr(), va.data_ptr(), kb.data_ptr(), vb.data_ptr(), n, end_bit,
                                               C.byref(in_b), None), "b200_sort_pairs_k32_static")
    else:
        sb = lib.b200_sort_scratch_bytes(n)
        scratch = torch.empty(sb, dtype=torch.uint8, device=DEV)
        L.check(lib.b200_sort_pairs_k32(ka.data_ptr(), va.data_ptr(), kb.data_ptr(), vb.data_ptr(), n, end_bit,
                                        scratch.data_ptr(), sb, C.byref(in_b), None), "b200_sort_pairs_k32")
    torch.cuda.synchronize()
    ks, vs = (kb, vb) if in_b.value else (ka, va)
    return ks.cpu(), vs.cpu(), in_b.value


def test_static_scratch_sort_equals_the_caller_scratch_sort_twice_in_a_row(pkg):
    lib, L = pkg._lib.lib, pkg._lib
    # two calls in a row sharing the library's array: a large one first, then a smaller one with fewer passes, whose
    # scratch region still holds the first call's histograms / status words unless the call clears them
    for n, end_bit, seed in ((1_000_003, 32, 1), (37_511, 20, 2), (1 << 20, 32, 3)):
        keys, k32, vals = _inputs(n, end_bit, seed)
        ks_s, vs_s, b_s = _run(lib, L, True, k32, vals, n, end_bit)
        ks_r, vs_r, b_r = _run(lib, L, False, k32, vals, n, end_bit)
        assert b_s == b_r
        assert torch.equal(ks_s, ks_r) and torch.equal(vs_s, vs_r)  # bit-exact
        order = torch.sort(keys, stable=True).indices  # and both equal a stable sort
        assert torch.equal(vs_s, vals[order])
        assert torch.equal(ks_s.to(torch.int64) & 0xFFFFFFFF, keys[order])


def test_static_scratch_sort_rejects_sizes_beyond_its_maximum(pkg):
    lib = pkg._lib.lib
    t = torch.zeros(8, dtype=torch.int32, device=DEV)
    rc = lib.b200_sort_pairs_k32_static(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), (1 << 20) + 1, 32, None, None)
    assert rc == -1 and b"B200_SORT_STATIC_MAX_KEYS" in lib.b200_last_error()
    in_b = C.c_int(-1)
    assert lib.b200_sort_pairs_k32_static(None, None, None, None, 0, 32, C.byref(in_b), None) == 0 and in_b.value == 0
