"""BatchNorm backward through the entry point that allocates nothing (``mpb_bn_bwd_noalloc``).

Some hosts of this library cannot hand us scratch memory: they have already given the whole GPU to their own
memory pool, or they call us from a place where cudaMalloc's implicit device synchronisation is unacceptable.
For them the library keeps the 2*C floats of reduction scratch in a fixed device array of its own.  This test
drives that entry point straight through the C ABI (ctypes, raw pointers) and checks, on two different calls in
a row -- the second one finds the first one's sums still in the shared array -- that it produces what the
ordinary ``mpb_bn_bwd`` (caller-provided scratch) produces, and that an over-sized C is refused.

The file name sorts last on purpose: with ``pytest -x`` every other test has already run if this one fails.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

c_void_p, c_int, c_ll = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong


def _lib():
    from mpi_pytorch_b200.build import lib_path
    return ctypes.CDLL(lib_path())


def _ptr(t):
    return c_void_p(0 if t is None else t.data_ptr())


def _case(seed, n, c, h, w, relu, with_res, dtype):
    g = torch.Generator(device="cuda").manual_seed(seed)
    shape = (n, h, w, c)                                       # NHWC, C contiguous
    x = (torch.randn(shape, device="cuda", generator=g) * 2 + 0.3).to(dtype)
    z = torch.randn(shape, device="cuda", generator=g).to(dtype)   # forward output (only its sign matters for ReLU)
    dz = torch.randn(shape, device="cuda", generator=g).to(dtype)
    gamma = torch.rand(c, device="cuda", generator=g) + 0.5
    xf = x.float().reshape(-1, c)
    mean = xf.mean(0)
    invstd = (xf.var(0, unbiased=False) + 1e-5).rsqrt()
    return x, z, dz, gamma, mean.contiguous(), invstd.contiguous(), relu, with_res


def _run(lib, fn_name, case, use_scratch):
    x, z, dz, gamma, mean, invstd, relu, with_res = case
    c = x.shape[-1]
    m = x.numel() // c
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if with_res else None
    dgamma = torch.empty(c, device="cuda")
    dbeta = torch.empty(c, device="cuda")
    dt = 1 if x.dtype == torch.bfloat16 else 0
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    fn = getattr(lib, fn_name)
    if use_scratch:
        dsum = torch.empty(2 * c, device="cuda")
        rc = fn(_ptr(dz), _ptr(z), _ptr(x), _ptr(gamma), _ptr(mean), _ptr(invstd), _ptr(dsum), _ptr(dx), _ptr(dres),
                _ptr(dgamma), _ptr(dbeta), c_ll(m), c_int(c), c_int(int(relu)), c_int(dt), stream)
    else:
        rc = fn(_ptr(dz), _ptr(z), _ptr(x), _ptr(gamma), _ptr(mean), _ptr(invstd), _ptr(dx), _ptr(dres),
                _ptr(dgamma), _ptr(dbeta), c_ll(m), c_int(c), c_int(int(relu)), c_int(dt), stream)
    torch.cuda.synchronize()
    assert rc == 0, "{} returned {}".format(fn_name, rc)
    return dx, dres, dgamma, dbeta


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    cases = [_case(11, 6, 128, 9, 9, True, True, torch.bfloat16),       # first call: sets the scratch address up
             _case(12, 4, 64, 14, 14, False, False, torch.float32)]     # second call: scratch still holds call 1's sums
    for case in cases:
        want = _run(lib, "mpb_bn_bwd", case, use_scratch=True)
        got = _run(lib, "mpb_bn_bwd_noalloc", case, use_scratch=False)
        tol = 3e-2 if case[0].dtype == torch.bfloat16 else 1e-4        # same bands as tests/test_kernels_gpu.py
        for a, b in zip(got, want):
            if a is None:
                assert b is None
                continue
            assert torch.allclose(a.float(), b.float(), atol=tol * 4, rtol=tol)
    # and against first principles for the fp32 case (plain PyTorch autograd on the same data)
    x, z, dz, gamma, mean, invstd, relu, _ = cases[1]
    xr = x.float().reshape(-1, x.shape[-1]).requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    y = (xr - xr.mean(0)) * (xr.var(0, unbiased=False) + 1e-5).rsqrt() * gr
    y.backward(dz.float().reshape(-1, x.shape[-1]))
    dx, _, dgamma, _ = _run(lib, "mpb_bn_bwd_noalloc", cases[1], use_scratch=False)
    assert torch.allclose(dx.float().reshape(-1, x.shape[-1]), xr.grad, atol=4e-4, rtol=1e-3)
    assert torch.allclose(dgamma, gr.grad, atol=3e-3, rtol=1e-3)


def test_noalloc_rejects_more_channels_than_its_scratch_holds():
    lib = _lib()
    t = torch.zeros(8, device="cuda")
    rc = lib.mpb_bn_bwd_noalloc(_ptr(t), _ptr(t), _ptr(t), _ptr(t), _ptr(t), _ptr(t), _ptr(t), c_void_p(0), _ptr(t),
                                _ptr(t), c_ll(1), c_int(4096), c_int(0), c_int(0), c_void_p(0))
    assert rc == -4
    # an empty batch is refused as well instead of launching zero-sized grids
    rc = lib.mpb_bn_bwd_noalloc(_ptr(t), _ptr(t), _ptr(t), _ptr(t), _ptr(t), _ptr(t), _ptr(t), c_void_p(0), _ptr(t),
                                _ptr(t), c_ll(0), c_int(8), c_int(0), c_int(0), c_void_p(0))
    assert rc == -1
