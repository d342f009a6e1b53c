"""CPU: host-side pieces of the rasterizer's launch path that need no GPU — the workspace carving, the argument holders that
must survive autocast's argument cast, the static-scratch entry point's size check."""
import collections.abc
import importlib

import torch


def test_carve_gives_disjoint_aligned_ranges(pkg):
    rz = importlib.import_module("3d-gaussian-splatting_b200.rasterization")
    sizes = (4 * 1000, 4 * 1000, 0, 7, 4 * 512, 300)
    owner, ptrs = rz._carve(torch.device("cpu"), sizes)
    base = owner.data_ptr()
    assert len(ptrs) == len(sizes) and ptrs[0] == base
    end = base + owner.numel()
    prev_end = base
    for p, b in zip(ptrs, sizes):
        assert (p - base) % 256 == 0 and p >= prev_end and p + b <= end
        prev_end = p + b
    owner0, ptrs0 = rz._carve(torch.device("cpu"), (0,))
    assert owner0.numel() >= 256 and ptrs0 == [owner0.data_ptr()]


def test_holders_are_not_rebuilt_by_custom_fwd_argument_cast(pkg):
    """torch.amp.custom_fwd casts the Function's arguments under autocast and REBUILDS Mapping / Iterable arguments; the
    holders the forward / backward write into must come through as the same objects."""
    from torch.amp.autocast_mode import _cast

    rz = importlib.import_module("3d-gaussian-splatting_b200.rasterization")
    st = importlib.import_module("3d-gaussian-splatting_b200.store")
    aux, state = rz._Aux(), st.GradState(pending_zero=True)
    for obj in (aux, state):
        assert not isinstance(obj, (collections.abc.Mapping, collections.abc.Iterable))
        assert _cast(obj, "cuda", torch.float32) is obj
    grad_into = (torch.zeros(2), torch.zeros(2), state)
    cast = _cast(grad_into, "cuda", torch.float32)
    assert cast[2] is state  # the tuple is rebuilt, its state element is not
    state["pending_zero"] = False
    assert cast[2].get("pending_zero") is False and state.get("missing") is None
    aux.radii = torch.zeros(1)
    assert aux["radii"] is aux.radii


def test_static_scratch_sort_rejects_oversize_without_a_gpu(pkg):
    lib = pkg._lib.lib
    rc = lib.b200_sort_pairs_k32_static(None, None, None, None, (1 << 20) + 1, 32, None, None)
    assert rc == -1 and b"B200_SORT_STATIC_MAX_KEYS" in lib.b200_last_error()
