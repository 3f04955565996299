"""ilfdt_colsum_static (include/ilfdt.h): the allocation-free variant of ilfdt_colsum, called through the C-ABI and
checked against the existing entry point. Named test_zz_* so pytest collects it LAST:
a failure here must not hide the parity tests. Passes alone in a fresh process and after the other tests (it makes
no assumption about what the shared static scratch holds when it starts).

The entry point has no counterpart in the reference: the comparison is
with ilfdt_colsum -- same kernels, same grid, so the sums must be BIT-identical, which is tighter than the 5e-5 the
colsum parity test (tests/test_parity_gpu.py::test_qmap_bias_cast_and_colsum) allows against the fp64 sum; that gate
is applied as well.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

COLSUM_TOL = 5e-5          # tests/test_parity_gpu.py::test_qmap_bias_cast_and_colsum


@pytest.fixture(scope="module")
def lib():
    from ilfdt import _lib
    return _lib.load()


def _dtype_code(t):
    from ilfdt import ops
    return ops._DT[t.dtype]


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _colsum_existing(lib, src):
    from ilfdt import _lib
    R, C = src.shape
    out = torch.empty(C, dtype=torch.float32, device=src.device)
    nws = lib.ilfdt_colsum_ws_bytes(R, C)
    ws = torch.empty(max(nws, 1), dtype=torch.uint8, device=src.device)
    _lib.check(lib.ilfdt_colsum(src.data_ptr(), _dtype_code(src), R, C, out.data_ptr(), ws.data_ptr(), nws, _stream()))
    return out


def _rel(a, ref):
    ref = ref.double()
    return ((a.double() - ref).abs().max() / ref.abs().max().clamp_min(1e-30)).item()


def test_colsum_static_matches_colsum_on_two_calls_in_a_row_and_allocates_nothing(lib):
    from ilfdt import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator().manual_seed(11)
    # two DIFFERENT inputs, the second one smaller (fewer CTAs, narrower rows): whatever the first call left in the
    # shared scratch beyond the second call's partial sums must not reach the second result
    first = (torch.randn(8192 * 3 + 5, 768, generator=g) + 0.25).to(dev)                       # q_map width, 592 CTAs
    second = (torch.randn(1003, 512, generator=g) - 0.5).to(torch.bfloat16).to(dev)           # 126 CTAs, bf16 rows
    want = [_colsum_existing(lib, first), _colsum_existing(lib, second)]
    outs = [torch.full((768,), float("nan"), device=dev), torch.full((512,), float("nan"), device=dev)]
    torch.cuda.synchronize()

    # a first call so that one-time set-up (symbol address, lazy module load) is not mistaken for an allocation
    _lib.check(lib.ilfdt_colsum_static(first.data_ptr(), _dtype_code(first), first.shape[0], 768, outs[0].data_ptr(), _stream()))
    torch.cuda.synchronize()
    outs[0].fill_(float("nan"))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    torch_alloc0, torch_reserved0 = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)

    # ---- the two calls in a row, same stream, nothing in between
    for src, out in zip((first, second), outs):
        _lib.check(lib.ilfdt_colsum_static(src.data_ptr(), _dtype_code(src), src.shape[0], src.shape[1], out.data_ptr(), _stream()))
    torch.cuda.synchronize()

    free1, _ = torch.cuda.mem_get_info(dev)
    assert torch.cuda.memory_allocated(dev) == torch_alloc0 and torch.cuda.memory_reserved(dev) == torch_reserved0
    assert free1 >= free0, "device memory in use grew by %d bytes across two ilfdt_colsum_static calls" % (free0 - free1)

    for src, out, ref in zip((first, second), outs, want):
        assert torch.isfinite(out).all()
        assert torch.equal(out, ref), "colsum_static differs from colsum by %.3e (same kernels, same grid: expected the same bits)" % _rel(out, ref)
        assert _rel(out, ref) <= COLSUM_TOL
        assert _rel(out, src.double().sum(0)) <= COLSUM_TOL

    # ---- and once more in the other order (the scratch now holds the SMALL call's partial sums)
    again = torch.full((768,), float("nan"), device=dev)
    _lib.check(lib.ilfdt_colsum_static(first.data_ptr(), _dtype_code(first), first.shape[0], 768, again.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert torch.equal(again, want[0])


def test_colsum_static_rejects_sizes_beyond_its_scratch(lib):
    from ilfdt import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    limit = 2 * 1024 * 1024                                   # ILFDT_COLSUM_STATIC_SCRATCH_BYTES
    R, C = 8192, 1024                                         # 592 CTAs x 1024 columns x 4 B = 2.3 MiB of partial sums
    assert lib.ilfdt_colsum_ws_bytes(R, C) > limit
    src = torch.ones(R, C, device=dev)
    out = torch.full((C,), -7.0, device=dev)
    rc = lib.ilfdt_colsum_static(src.data_ptr(), _dtype_code(src), R, C, out.data_ptr(), _stream())
    assert rc == -1                                            # ILFDT_E_INVALID_ARG
    msg = lib.ilfdt_last_error().decode()
    assert "static scratch" in msg and "ilfdt_colsum" in msg and str(limit) in msg, msg
    with pytest.raises(_lib.IlfdtError):
        _lib.check(rc)
    torch.cuda.synchronize()
    assert bool((out == -7.0).all()), "a rejected call must not touch the output"
    # the largest width still fits while the CTA count is small enough
    small = torch.ones(4096, C, device=dev)
    assert lib.ilfdt_colsum_ws_bytes(4096, C) <= limit
    _lib.check(lib.ilfdt_colsum_static(small.data_ptr(), _dtype_code(small), 4096, C, out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert torch.equal(out, torch.full((C,), 4096.0, device=dev))
