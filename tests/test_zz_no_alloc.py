"""soda_multibox_detection_static: the allocation-free variant of soda_multibox_detection (scratch in a file-scope
device array of the library, include/soda_b200.h).  Called straight through the C ABI and compared with the existing
entry point on the same inputs: same kernel, same arithmetic, so the rows must be bit-identical (the tolerance of the
other multibox tests, exact class / confidence and 1e-5 on the corners, is implied).

Collected last on purpose (file name).  Passes alone in a fresh process and after
the other tests.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _lib():
    from snn_for_object_detection_b200 import _lib

    return _lib


def _inputs(B, A, C, seed):
    g = torch.Generator().manual_seed(seed)
    ctr = torch.rand(A, 2, generator=g)
    wh = torch.rand(A, 2, generator=g) * 0.2 + 0.02
    anchors = torch.cat([ctr - wh / 2, ctr + wh / 2], dim=1)
    probs = torch.softmax(torch.randn(B, A, C, generator=g) * 2.0, dim=2)
    probs[:, ::7, 0] = 0.999
    offsets = torch.randn(B, A, 4, generator=g) * 0.5
    return probs.to(DEV).contiguous(), offsets.to(DEV).contiguous(), anchors.to(DEV).contiguous()


def _both(B, A, C, seed):
    """(rows of the existing entry point with a caller workspace, rows of the static-scratch entry point)."""
    m = _lib()
    lib = m.lib
    probs, offsets, anchors = _inputs(B, A, C, seed)
    stream = torch.cuda.current_stream().cuda_stream
    ws = torch.empty(lib.soda_multibox_workspace_bytes(B, A), dtype=torch.uint8, device=DEV)
    ref = torch.full((B, A, 6), float("nan"), device=DEV)
    m.check(lib.soda_multibox_detection(probs.data_ptr(), offsets.data_ptr(), anchors.data_ptr(), B, A, C, 0.1, 0.009999999,
                                        ws.data_ptr(), ref.data_ptr(), stream), "soda_multibox_detection")
    out = torch.full((B, A, 6), float("nan"), device=DEV)   # inputs and outputs are still the caller's
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    m.check(lib.soda_multibox_detection_static(probs.data_ptr(), offsets.data_ptr(), anchors.data_ptr(), B, A, C, 0.1,
                                               0.009999999, out.data_ptr(), stream), "soda_multibox_detection_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before           # nothing went through torch's allocator either
    return ref.cpu(), out.cpu()


def test_static_scratch_multibox_equals_the_workspace_entry_point_twice_in_a_row():
    # first call: GEN1 anchor count (keys sorted in shared memory); second: a smaller, different problem on top of the
    # first call's leftovers in the shared array; third: the first again
    for B, A, C, seed in [(5, 13545, 3, 1), (2, 4097, 4, 2), (5, 13545, 3, 1)]:
        ref, out = _both(B, A, C, seed)
        assert not torch.isnan(out).any()
        assert int((ref[..., 0] >= 0).sum()) > 10
        assert torch.equal(out[..., :2], ref[..., :2])
        assert torch.allclose(out[..., 2:], ref[..., 2:], atol=1e-5, rtol=1e-5)
        assert torch.equal(out, ref)


def test_static_scratch_large_anchor_set_uses_the_array_for_its_keys():
    # 40 000 anchors > 16 384: sort keys, flags and kept marks live in the scratch too
    for seed in (3, 4):
        ref, out = _both(1, 40000, 3, seed)
        assert torch.equal(out, ref)


def test_static_scratch_rejects_sizes_beyond_the_array():
    m = _lib()
    lib = m.lib
    cap = lib.soda_multibox_static_max_bytes()
    assert cap == 8 << 20
    B, A, C = 15, 13545, 3                                    # 14 GEN1 samples fit, 15 do not
    assert lib.soda_multibox_workspace_bytes(14, A) <= cap < lib.soda_multibox_workspace_bytes(B, A)
    probs, offsets, anchors = _inputs(B, A, C, 5)
    out = torch.zeros((B, A, 6), device=DEV)
    rc = lib.soda_multibox_detection_static(probs.data_ptr(), offsets.data_ptr(), anchors.data_ptr(), B, A, C, 0.1, 0.01,
                                            out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 2                                            # SODA_ERR_UNSUPPORTED
    msg = lib.soda_last_error().decode()
    assert "scratch bytes" in msg and str(cap) in msg
    with pytest.raises(m.SodaError):
        m.check(rc, "soda_multibox_detection_static")
    assert float(out.abs().max()) == 0                        # nothing was launched
