"""Error behaviour of the C-ABI boundary (include/ilfdt.h) on a real device, called through ctypes exactly as a foreign
binding would: every entry point returns an int32 status, 0 = ok, negative = ILFDT_E_*, and `ilfdt_last_error()` holds a
message that names what was wrong. A refused call must not touch its outputs, must leave no sticky CUDA error behind,
and the next valid call must work. The Python host side turns a non-zero status into `IlfdtError` -- nothing falls back
to PyTorch math.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

OK, E_INVALID_ARG, E_UNSUPPORTED_SHAPE, E_CUDA, E_NCCL, E_ARCH = 0, -1, -2, -3, -4, -5
F32, BF16 = 0, 1


@pytest.fixture(scope="module")
def lib():
    from ilfdt import _lib
    return _lib.load()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _msg(lib):
    m = lib.ilfdt_last_error()
    return m.decode() if m else ""


def _dev():
    return torch.device("cuda:0")


def _healthy(lib):
    """A valid call right after a refused one: works, and CUDA carries no sticky error."""
    x = torch.arange(8, dtype=torch.float32, device=_dev())
    out = torch.empty(8, dtype=torch.bfloat16, device=_dev())
    assert lib.ilfdt_cast_bf16(x.data_ptr(), out.data_ptr(), 8, _stream()) == OK
    torch.cuda.synchronize()
    assert torch.equal(out.float(), x)


def test_library_identity_and_device_check(lib):
    assert lib.ilfdt_abi_version() == 4
    assert lib.ilfdt_device_check(0) == OK
    rc = lib.ilfdt_device_check(torch.cuda.device_count() + 7)       # no such device
    assert rc in (E_CUDA, E_ARCH, E_INVALID_ARG) and _msg(lib) != ""
    _healthy(lib)


def test_tune_keys_are_validated(lib):
    assert lib.ilfdt_tune_set(-1, 1) == E_INVALID_ARG
    assert lib.ilfdt_tune_set(10_000, 1) == E_INVALID_ARG
    assert "tune" in _msg(lib).lower()
    assert lib.ilfdt_tune_get(10_000) == 0
    from ilfdt import _lib
    for name, key in _lib.TUNE.items():                              # every documented key round-trips
        old = lib.ilfdt_tune_get(key)
        assert lib.ilfdt_tune_set(key, old) == OK, name


def test_dtype_codes_are_validated(lib):
    a = torch.zeros(128, 64, dtype=torch.bfloat16, device=_dev())
    out = torch.full((128, 128), -3.0, device=_dev())
    rc = lib.ilfdt_gemm_nt(a.data_ptr(), a.data_ptr(), 7, 128, 128, 64, ctypes.c_float(1.0), 0, out.data_ptr(), 128, _stream())
    assert rc == E_INVALID_ARG and "dtype" in _msg(lib).lower()
    torch.cuda.synchronize()
    assert bool((out == -3.0).all())
    _healthy(lib)


def test_null_pointers_and_bad_shapes_are_refused_without_touching_outputs(lib):
    dev = _dev()
    z = torch.randn(4, 64, device=dev)
    w = torch.full((4, 64), -5.0, device=dev)
    # null input
    assert lib.ilfdt_sparsemax_fwd(0, 4, 64, w.data_ptr(), 0, 0, 0, _stream()) != OK
    # K = 0
    assert lib.ilfdt_sparsemax_fwd(z.data_ptr(), 4, 0, w.data_ptr(), 0, 0, 0, _stream()) != OK
    # a row that does not fit one CTA's shared memory (documented limit: K <= 51200)
    rc = lib.ilfdt_sparsemax_fwd(z.data_ptr(), 1, 60000, w.data_ptr(), 0, 0, 0, _stream())
    assert rc == E_UNSUPPORTED_SHAPE and "51200" in _msg(lib)
    torch.cuda.synchronize()
    assert bool((w == -5.0).all())
    # gemm: leading dimension smaller than the row length
    a = torch.zeros(128, 64, dtype=torch.bfloat16, device=dev)
    out = torch.full((128, 128), -3.0, device=dev)
    rc = lib.ilfdt_gemm_nt(a.data_ptr(), a.data_ptr(), BF16, 128, 128, 64, ctypes.c_float(1.0), 0, out.data_ptr(), 64, _stream())
    assert rc == E_INVALID_ARG and "gemm_nt" in _msg(lib)
    rc = lib.ilfdt_gemm_nt(0, a.data_ptr(), BF16, 128, 128, 64, ctypes.c_float(1.0), 0, out.data_ptr(), 128, _stream())
    assert rc == E_INVALID_ARG and "null" in _msg(lib).lower()
    torch.cuda.synchronize()
    assert bool((out == -3.0).all())
    _healthy(lib)


def test_empty_problems_are_ok_and_launch_nothing(lib):
    dev = _dev()
    out = torch.full((4,), -1.0, device=dev)
    assert lib.ilfdt_cast_bf16(0, 0, 0, _stream()) == OK
    assert lib.ilfdt_sparsemax_fwd(0, 0, 64, 0, 0, 0, 0, _stream()) == OK
    assert lib.ilfdt_gemm_nt(0, 0, BF16, 0, 128, 64, ctypes.c_float(1.0), 0, 0, 128, _stream()) == OK
    assert lib.ilfdt_copy(0, 0, 0, _stream()) == OK
    torch.cuda.synchronize()
    assert bool((out == -1.0).all())


def test_workspace_too_small_is_an_error_not_a_fault(lib):
    dev = _dev()
    b, B, D = 64, 64, 64
    x = torch.randn(b, D, device=dev).to(torch.bfloat16)
    scale = torch.tensor([10.0], device=dev)
    lse = torch.full((2, b), -9.0, device=dev)
    diag = torch.full((2, b), -9.0, device=dev)
    loss = torch.full((1,), -9.0, device=dev)
    need = lib.ilfdt_nce_ws_bytes(BF16, b, B, D)
    assert need > 0
    ws = torch.zeros(16, dtype=torch.uint8, device=dev)
    rc = lib.ilfdt_nce_fwd(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), BF16, b, B, D, scale.data_ptr(), 0,
                           lse[0].data_ptr(), lse[1].data_ptr(), diag[0].data_ptr(), diag[1].data_ptr(), loss.data_ptr(),
                           ws.data_ptr(), ws.numel(), _stream())
    assert rc == E_INVALID_ARG and "workspace" in _msg(lib).lower()
    rc = lib.ilfdt_nce_fwd(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), BF16, b, B, D, scale.data_ptr(), 0,
                           lse[0].data_ptr(), lse[1].data_ptr(), diag[0].data_ptr(), diag[1].data_ptr(), loss.data_ptr(),
                           0, need, _stream())
    assert rc == E_INVALID_ARG
    torch.cuda.synchronize()
    assert bool((loss == -9.0).all()) and bool((lse == -9.0).all())
    # with the workspace it asks for, the same call runs
    ws = torch.zeros(need, dtype=torch.uint8, device=dev)
    rc = lib.ilfdt_nce_fwd(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), BF16, b, B, D, scale.data_ptr(), 0,
                           lse[0].data_ptr(), lse[1].data_ptr(), diag[0].data_ptr(), diag[1].data_ptr(), loss.data_ptr(),
                           ws.data_ptr(), need, _stream())
    assert rc == OK
    torch.cuda.synchronize()
    assert torch.isfinite(loss).all() and abs(float(loss)) < 1e-3     # identical image / text rows: the diagonal dominates
    assert bool((lse != -9.0).all())


def test_python_host_side_raises_instead_of_falling_back(lib):
    from ilfdt import _lib, ops
    dev = _dev()
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        ops.cast_bf16(torch.zeros(4, 4))                              # CPU tensor: there is no CPU path
    with pytest.raises(RuntimeError, match="contiguous"):
        ops.cast_bf16(torch.zeros(8, 8, device=dev).t())
    with pytest.raises(_lib.IlfdtError, match="51200"):
        ops.sparsemax_fwd(torch.zeros(1, 60000, device=dev))
    # and the refusal left the library usable
    w = ops.sparsemax_fwd(torch.zeros(2, 8, device=dev))[0]
    assert torch.allclose(w, torch.full((2, 8), 0.125, device=dev))
