"""vu_sort_pairs_desc_f32_static (allocation-free variant: scratch in a file-scope device array) against the existing
vu_sort_pairs_desc_f32 with a caller workspace, through the C ABI, on two calls in a row (the second call runs over the first
call's stale scratch). Integer work: the bar is bit-exact. Collected last on purpose (tests/test_zz_*)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _sort_both(lib, score, vals, stream):
    n = score.numel()
    dev = score.device
    ws_bytes = lib.vu_sort_workspace_bytes(n)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    k_ref, v_ref = torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.int32, device=dev)
    k_new, v_new = torch.full((n,), -1, dtype=torch.int32, device=dev), torch.full((n,), -1, dtype=torch.int32, device=dev)
    rc = lib.vu_sort_pairs_desc_f32(score.data_ptr(), vals.data_ptr(), n, k_ref.data_ptr(), v_ref.data_ptr(), ws.data_ptr(),
                                    ws_bytes, stream)
    assert rc == 0, lib.vu_last_error().decode()
    torch.cuda.synchronize()
    # everything the new entry point touches is allocated above: the call itself must not allocate
    before = torch.cuda.memory_allocated(dev)
    rc = lib.vu_sort_pairs_desc_f32_static(score.data_ptr(), vals.data_ptr(), n, k_new.data_ptr(), v_new.data_ptr(), stream)
    assert rc == 0, lib.vu_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before
    return (k_ref, v_ref), (k_new, v_new)


def test_static_scratch_sort_matches_workspace_sort(cuda_dev):
    from vaeunet_b200._lib import load
    lib = load()
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=cuda_dev).manual_seed(77)
    # two calls in a row with different sizes and contents; ties (quantised scores), +-0, inf
    for n, quant in ((262144, 64.0), (100003, 0.0)):
        score = torch.randn(n, device=cuda_dev, generator=g)
        if quant:
            score = torch.round(score * quant) / quant
        score[:4] = torch.tensor([0.0, -0.0, float("inf"), -float("inf")], device=cuda_dev)
        vals = torch.arange(n, dtype=torch.int32, device=cuda_dev)
        (k_ref, v_ref), (k_new, v_new) = _sort_both(lib, score, vals, stream)
        assert torch.equal(k_new, k_ref) and torch.equal(v_new, v_ref)
        s_sorted = score[v_new.long()]
        assert bool((s_sorted[:-1] >= s_sorted[1:]).all())                       # descending
        assert torch.equal(torch.sort(v_new.long()).values, torch.arange(n, device=cuda_dev))   # a permutation


def test_static_scratch_sort_rejects_sizes_beyond_its_limit(cuda_dev):
    from vaeunet_b200._lib import load
    lib = load()
    n = 262144 + 1
    score = torch.zeros(n, device=cuda_dev)
    vals = torch.zeros(n, dtype=torch.int32, device=cuda_dev)
    out = torch.zeros(n, dtype=torch.int32, device=cuda_dev)
    rc = lib.vu_sort_pairs_desc_f32_static(score.data_ptr(), vals.data_ptr(), n, out.data_ptr(), out.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b"static scratch limit" in lib.vu_last_error()
