"""GPU: the tcgen05 GEMM of the bf16 training engine (csrc/gemm_tc.cuh) through the C ABI against torch float32 matmul of the
same bf16 inputs.  fp32 accumulation of exact bf16 products: 1e-4 relative (max|a-b| / max|b|) for float32 outputs, 1e-2 for
bf16 outputs.  Shapes cover K and M / N tails, several N tiles, the persistent tile loop, views with a leading dimension, bias /
ReLU / accumulate, and the weight-gradient form (reduction over the rows, both operands MN-major, split over CTAs)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize('M,N,K,lda,ldb,ldo', [(300, 72, 200, 208, 200, 80), (128, 16, 64, 64, 64, 16), (5000, 400, 147, 152, 152, 400),
                                               (20000, 96, 1152, 1152, 1152, 480), (777, 8, 1024, 1024, 1024, 8)])
def test_gemm_k_major(M, N, K, lda, ldb, ldo):
    import torch
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device='cuda').manual_seed(M + N + K)
    A = torch.randn((M, lda), device='cuda', generator=g).bfloat16()
    B = torch.randn((N, ldb), device='cuda', generator=g).bfloat16()
    bias = torch.randn((N,), device='cuda', generator=g)
    ref = A[:, :K].float() @ B[:, :K].float().t()
    s = torch.cuda.current_stream().cuda_stream
    # float32 output, no epilogue
    out = torch.full((M, ldo), 7.0, device='cuda')
    _lib.check(L.fcnod_gemm_bf16_device(0, A.data_ptr(), lda, B.data_ptr(), ldb, M, N, K, out.data_ptr(), ldo, 1, None, 0, 0, s))
    torch.cuda.synchronize()
    assert _rel(out[:, :N], ref) < 1e-4
    assert bool((out[:, N:] == 7.0).all())                     # columns beyond N untouched
    # bf16 output with bias + ReLU, then accumulate on top of it
    outb = torch.zeros((M, ldo), device='cuda', dtype=torch.bfloat16)
    _lib.check(L.fcnod_gemm_bf16_device(0, A.data_ptr(), lda, B.data_ptr(), ldb, M, N, K, outb.data_ptr(), ldo, 0, bias.data_ptr(), 1, 0, s))
    torch.cuda.synchronize()
    want = torch.relu(ref + bias)
    assert _rel(outb[:, :N].float(), want) < 1e-2
    assert bool((outb[:, N:] == 0).all())                      # the TMA store clips at N
    first = outb.clone()
    _lib.check(L.fcnod_gemm_bf16_device(0, A.data_ptr(), lda, B.data_ptr(), ldb, M, N, K, outb.data_ptr(), ldo, 0, None, 0, 1, s))
    torch.cuda.synchronize()
    assert _rel(outb[:, :N].float(), first[:, :N].float() + ref) < 1.5e-2
    assert bool((outb[:, N:] == 0).all())


@pytest.mark.parametrize('PX,M,N,lda,ldb,ldo', [(1000, 96, 147, 480, 152, 128), (64, 128, 64, 128, 64, 128), (25088, 208, 864, 512, 864, 256),
                                                (3000, 16, 400, 16, 400, 64), (4096, 384, 192, 1024, 832, 384)])
def test_gemm_mn_major_weight_gradient_form(PX, M, N, lda, ldb, ldo):
    import torch
    from fcn_object_detector_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device='cuda').manual_seed(PX + M + N)
    A = torch.randn((PX, lda), device='cuda', generator=g).bfloat16()       # dY view: M of lda channels
    B = torch.randn((PX, ldb), device='cuda', generator=g).bfloat16()       # X / im2col view
    ref = B[:, :N].float().t() @ A[:, :M].float()                            # [N][M]
    out = torch.full((N, ldo), 0.5, device='cuda')
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(L.fcnod_gemm_bf16_device(1, A.data_ptr(), lda, B.data_ptr(), ldb, M, N, PX, out.data_ptr(), ldo, 1, None, 0, 0, s))
    torch.cuda.synchronize()
    assert _rel(out[:, :M] - 0.5, ref) < 1e-4                                # accumulated ON TOP of what was there
    assert bool((out[:, M:] == 0.5).all())
