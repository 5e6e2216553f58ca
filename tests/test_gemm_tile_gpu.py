"""GPU: the tensor-core GEMM tile primitive (operand layout, UMMA descriptors, TMEM readback)
in isolation, through the C ABI (sgnn_pack_weight + sgnn_debug_gemm)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(k, n, tc):
    from sgnn_b200 import _lib as L
    g = torch.Generator(device="cpu").manual_seed(k * 1000 + n)
    a = torch.randn(128, k, generator=g).cuda()
    w = torch.randn(n, k, generator=g).cuda()
    img = torch.zeros(n * k, device="cuda")
    d = torch.zeros(128, n, device="cuda")
    L.check(L.lib().sgnn_pack_weight(L.ptr(w), k, 0, n, 0, k, k, C.c_float(1.0), int(tc), L.ptr(img), L.stream_ptr()))
    L.check(L.lib().sgnn_debug_gemm(L.ptr(a), L.ptr(img), k, n, L.ptr(d), int(tc), L.stream_ptr()))
    torch.cuda.synchronize()
    return a, w, d


@pytest.mark.parametrize("k,n", [(32, 32), (64, 64), (128, 64), (64, 128), (256, 64)])
def test_gemm_tile_fp32(k, n):
    a, w, d = _run(k, n, False)
    ref = (a.double() @ w.double().T).float()
    assert torch.allclose(d, ref, atol=2e-5 * k ** 0.5, rtol=1e-5)


@pytest.mark.parametrize("k,n", [(32, 32), (64, 64), (128, 64), (64, 128), (256, 64)])
def test_gemm_tile_tf32(k, n):
    a, w, d = _run(k, n, True)
    ref = (a.double() @ w.double().T).float()
    # operands rounded to TF32 (rel 2^-11 each), fp32 accumulation
    assert torch.allclose(d, ref, atol=2e-3 * k ** 0.5, rtol=2e-3)
    # and it really is the TF32 product: compare with operands rounded the same way
    def rt(x):
        return ((x.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)
    ref_t = (rt(a).double() @ rt(w).double().T).float()
    assert torch.allclose(d, ref_t, atol=3e-5 * k ** 0.5, rtol=1e-5)


@pytest.mark.parametrize("k,n", [(32, 32), (64, 64), (128, 64), (64, 128), (128, 128)])
def test_gemm_tile_tf32_a_in_tmem(k, n):
    """A operand staged in tensor memory (tcgen05.st + tcgen05.mma with A from TMEM)."""
    a, w, d = _run(k, n, 2)
    def rt(x):
        return ((x.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)
    ref_t = (rt(a).double() @ rt(w).double().T).float()
    assert torch.allclose(d, ref_t, atol=3e-5 * k ** 0.5, rtol=1e-5)
