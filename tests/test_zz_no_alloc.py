"""b200sd_colsum_static — the allocation-free variant of b200sd_colsum (include/b200sd.h): scratch in a static __device__
array of the library, address fetched with cudaGetSymbolAddress. Collected last on purpose.

Checked through the raw C-ABI against the existing entry point (same kernel and row blocks -> bit-identical), twice in
a row (the shared scratch keeps the previous call's contents), for both output forms, plus the size limit."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _raw(rt):
    return rt.lib(), rt.stream_ptr()


def _existing(rt, x, n, hw, c, per_image):
    """b200sd_colsum with caller-owned scratch, called directly (part + zeroed counters from torch)."""
    L, st = _raw(rt)
    pb = int(L.b200sd_colsum_blocks(n, hw, c))
    part = torch.empty(n * pb * c, device=x.device, dtype=torch.float32)
    counter = torch.zeros(64, device=x.device, dtype=torch.int32)
    out = torch.full((n, c) if per_image else (c,), float("nan"), device=x.device, dtype=torch.float32)
    rc = L.b200sd_colsum(x.data_ptr(), x.stride(0), n, hw, c, part.data_ptr(), out.data_ptr(), int(per_image),
                         counter.data_ptr(), st)
    assert rc == 0, L.b200sd_last_error().decode()
    return out


def _static(rt, x, n, hw, c, per_image, out):
    L, st = _raw(rt)
    rc = L.b200sd_colsum_static(x.data_ptr(), x.stride(0), n, hw, c, out.data_ptr(), int(per_image), st)
    assert rc == 0, L.b200sd_last_error().decode()
    return out


@pytest.mark.parametrize("n,hw,c,per_image", [(4, 1024, 384, False), (4, 1024, 384, True), (3, 77, 136, False),
                                              (16, 256, 1152, True)])
def test_static_scratch_matches_caller_scratch_twice(rt, n, hw, c, per_image):
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(n * 1000 + c)
    outs = []
    for call in range(2):   # two calls in a row: different data, the scratch still holds the first call's partials
        x = torch.randn(n * hw, c, generator=g).to(torch.bfloat16).to(dev)
        ref = _existing(rt, x, n, hw, c, per_image)
        out = torch.full_like(ref, float("nan"))
        before = torch.cuda.memory_allocated(dev)
        _static(rt, x, n, hw, c, per_image, out)
        assert torch.cuda.memory_allocated(dev) == before
        torch.cuda.synchronize()
        assert torch.equal(out, ref), f"call {call}: max diff {(out - ref).abs().max().item()}"
        # and against plain fp32 torch on the same bf16 inputs (the tolerance the colsum test uses)
        want = x.float().view(n, hw, c).sum(1) if per_image else x.float().sum(0)
        torch.testing.assert_close(out, want, rtol=2e-4, atol=2e-3 * (hw * (1 if per_image else n)) ** 0.5)
        outs.append(out.clone())
    assert not torch.equal(outs[0], outs[1])


def test_static_scratch_capturable_and_allocation_free(rt):
    """Replays inside a CUDA graph (set-up call made outside capture first), result identical to the eager call."""
    dev = torch.device("cuda:0")
    n, hw, c = 2, 512, 256
    x = torch.randn(n * hw, c, generator=torch.Generator().manual_seed(5)).to(torch.bfloat16).to(dev)
    out = torch.zeros(c, device=dev, dtype=torch.float32)
    ref = _existing(rt, x, n, hw, c, False)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        _static(rt, x, n, hw, c, False, out)      # set-up (cudaGetSymbolAddress) outside capture
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    out.zero_()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        _static(rt, x, n, hw, c, False, out)
    graph.replay()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


def test_static_scratch_rejects_oversize(rt):
    L, st = _raw(rt)
    L.b200sd_colsum_static_max_floats.restype = C.c_int64
    limit = int(L.b200sd_colsum_static_max_floats())
    assert limit == 512 * 1024
    dev = torch.device("cuda:0")
    n, hw, c = 256, 64, 4096          # 256 * 1 * 4096 = 1 Mi floats of partials > limit
    assert n * int(L.b200sd_colsum_blocks(n, hw, c)) * c > limit
    x = torch.zeros(n * hw, c, device=dev, dtype=torch.bfloat16)
    out = torch.zeros(c, device=dev, dtype=torch.float32)
    rc = L.b200sd_colsum_static(x.data_ptr(), x.stride(0), n, hw, c, out.data_ptr(), 0, st)
    assert rc == -1
    msg = L.b200sd_last_error().decode()
    assert "exceed the static scratch" in msg and "b200sd_colsum" in msg
