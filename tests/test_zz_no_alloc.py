"""gsb_dice_sums_noalloc: the allocation-free variant of gsb_dice_sums (scratch in a file-scope `static __device__` array
of the library, include/genseg_b200.h).  Called through the C ABI and compared with the existing entry point on two calls
in a row (the array keeps the previous call's contents).  Collected last on purpose.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from semantic_segmentation_b200._lib import load  # noqa: E402

DEV = "cuda"


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dice_sums_workspace(lib, a, b, slices, inner, rows):
    """The existing entry point with a caller workspace of exactly `rows` partial rows."""
    sums = torch.empty(slices * 3, device=DEV, dtype=torch.float64)
    ws = torch.empty(rows * slices * 3, device=DEV, dtype=torch.float64)
    rc = lib.gsb_dice_sums(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), slices, inner, C.c_void_p(sums.data_ptr()),
                           C.c_void_p(ws.data_ptr()), ws.numel() * 8, _stream())
    assert rc == 0, lib.gsb_last_error().decode()
    return sums


def _dice_sums_noalloc(lib, a, b, slices, inner, sums):
    return lib.gsb_dice_sums_noalloc(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), slices, inner,
                                     C.c_void_p(sums.data_ptr()), _stream())


@pytest.mark.parametrize("slices,inner", [(4, 256 * 256), (64, 70001), (1, 5)])
def test_dice_sums_noalloc_matches_workspace_form(slices, inner):
    lib = load()
    assert int(lib.gsb_dice_sums_noalloc_max_slices()) == 64
    g = torch.Generator().manual_seed(77)
    outs = []
    sums = torch.empty(slices * 3, device=DEV, dtype=torch.float64)        # caller-owned output, allocated before the calls
    for call in range(2):                                                 # two calls in a row, different data
        a = torch.rand(slices, inner, generator=g).to(DEV)
        b = (torch.rand(slices, inner, generator=g) > 0.5).float().to(DEV)
        rows = min((inner + 4095) // 4096, 1024, 12288 // (3 * slices))   # the row count the no-alloc form uses
        want = _dice_sums_workspace(lib, a, b, slices, inner, max(rows, 1))
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = _dice_sums_noalloc(lib, a, b, slices, inner, sums)
        assert rc == 0, lib.gsb_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before                    # nothing taken from torch's allocator
        assert torch.equal(sums, want), (call, (sums - want).abs().max().item())   # same rows, same order: bit-identical
        ref = torch.stack([(a.double() * b.double()).sum(1), a.double().sum(1), b.double().sum(1)], 1).reshape(-1)
        assert ((sums - ref).abs() / ref.abs().clamp_min(1.0)).max().item() < 1e-6  # 1e-6: the bound tests/test_gpu_unet.py holds on the Dice values
        outs.append(sums.clone())
    assert not torch.equal(outs[0], outs[1])                              # the second call did not return stale rows


def test_dice_sums_noalloc_rejects_oversize():
    lib = load()
    a = torch.zeros(65, 8, device=DEV)
    sums = torch.empty(65 * 3, device=DEV, dtype=torch.float64)
    rc = _dice_sums_noalloc(lib, a, a, 65, 8, sums)
    assert rc != 0
    msg = lib.gsb_last_error().decode()
    assert "65 slices exceed" in msg and "max 64" in msg
