"""tcgen05 / TMA / TMEM GEMM vs a plain PyTorch fp32 reference of the same (bf16-rounded) operands."""
import pytest
import torch

from pytorch_distributed_rnn_b200.ops.gemm import cast_bf16, gemm_bf16, gemm_status

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 128, 256), (256, 384, 512), (1000, 520, 264), (64, 4096, 1024),
                                   (5, 24, 8), (4096, 128, 32),
                                   (4096, 1024, 128), (3000, 1000, 72),    # large enough for the 128x256 tiles
                                   (4096, 2560, 192), (2176, 4356, 200)])  # persistent kernel: several tiles per CTA, tails
def test_gemm_matches_fp32_reference(M, N, K):
    torch.manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    want = a.float() @ b.float().t() + bias
    got = gemm_bf16(a, b, bias)
    torch.cuda.synchronize()
    assert gemm_status() == 0
    err = float((got - want).abs().max())
    assert err <= 1e-3 * (K ** 0.5) + 1e-3, err
    # accumulate into fp32 C, and bf16 output
    got2 = gemm_bf16(a, b, None, out=got.clone(), accumulate=True)
    assert torch.allclose(got2, want + (a.float() @ b.float().t()), atol=2e-3 * (K ** 0.5) + 1e-2, rtol=1e-3)
    got3 = gemm_bf16(a, b, bias, out_dtype=torch.bfloat16)
    assert torch.allclose(got3.float(), want, atol=0.05 * (K ** 0.5), rtol=2e-2)
    assert gemm_status() == 0


def test_cast_and_transpose_kernels():
    x = torch.randn(77, 130, device="cuda")
    assert torch.equal(cast_bf16(x), x.bfloat16())
    assert torch.equal(cast_bf16(x, transpose=True), x.t().contiguous().bfloat16())
