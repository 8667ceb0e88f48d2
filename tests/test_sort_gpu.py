"""gsm_sort_pairs32 (the hand-written replacement of cub::DeviceRadixSort::SortPairs at
simple_knn.cu:213-214) against a stable reference sort: bit-exact keys AND values (stability),
empty / single / ragged / multi-tile sizes, partial key ranges (end_bit), all-equal keys."""
import ctypes

import pytest
import torch

from gaussmart_b200 import _lib

pytestmark = pytest.mark.gpu


def run_sort(keys, vals, end_bit):
    lib = _lib.load()
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    tb = lib.gsm_sort_pairs32_temp_bytes(n)
    tmp = torch.empty(max(tb, 1), dtype=torch.uint8, device=keys.device)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    _lib.check(lib.gsm_sort_pairs32(p(keys), p(ko), p(vals), p(vo), n, end_bit, p(tmp), tb,
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "sort_pairs32")
    torch.cuda.synchronize()
    return ko, vo


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 4095, 4096, 4097, 100_003, 5_000_000])
@pytest.mark.parametrize("end_bit,key_bits", [(30, 30), (32, 32), (13, 32), (8, 8), (30, 4)])
def test_sort_pairs32_is_a_stable_sort(n, end_bit, key_bits):
    g = torch.Generator(device="cpu").manual_seed(1234 + n + end_bit)
    keys = torch.randint(0, 2 ** key_bits, (n,), generator=g, dtype=torch.int64).to(torch.int32 if key_bits < 32 else torch.int64)
    keys = (keys & 0xffffffff).to(torch.int64)
    keys_u = keys.cuda()
    k32 = torch.where(keys_u >= 2 ** 31, keys_u - 2 ** 32, keys_u).to(torch.int32)   # same bits as uint32
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = run_sort(k32, vals, end_bit)
    # reference: stable sort on the masked key; values are the original positions => ties ascending
    mask = (1 << end_bit) - 1
    order = torch.sort(keys_u & mask, stable=True).indices
    assert torch.equal(vo.long(), order)
    assert torch.equal(ko, k32[order])


def test_sort_pairs32_all_equal_and_presorted():
    n = 70_001
    k = torch.full((n,), 12345, dtype=torch.int32, device="cuda")
    v = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = run_sort(k, v, 30)
    assert torch.equal(ko, k) and torch.equal(vo, v)
    k = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = run_sort(k, v.flip(0).contiguous(), 30)
    assert torch.equal(ko, k) and torch.equal(vo, v.flip(0))


def test_sort_pairs32_rejects_small_temp():
    lib = _lib.load()
    k = torch.zeros(100, dtype=torch.int32, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    assert lib.gsm_sort_pairs32(p(k), p(k.clone()), p(k), p(k.clone()), 100, 30, None, 0, None) == _lib.GSM_ERR_WORKSPACE_TOO_SMALL
    assert lib.gsm_sort_pairs32(p(k), p(k.clone()), p(k), p(k.clone()), 100, 40, None, 0, None) == _lib.GSM_ERR_INVALID_ARGUMENT
