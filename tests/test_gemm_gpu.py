"""tcgen05 GEMM building block vs a plain fp32 PyTorch matmul of the same bf16 inputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, bias, relu):
    c = a.float() @ b.float().t()
    if bias is not None:
        c = c + bias
    return torch.relu(c) if relu else c


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (128, 256, 128), (256, 128, 256), (2048, 256, 1248),
                                   (300, 48, 200), (2048, 16, 72), (1000, 400, 2496)])
def test_gemm_k_major(M, N, K):
    from deepfm_b200.ops.gemm import gemm_bf16
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    Kp = (K + 7) // 8 * 8
    a = torch.randn(M, Kp, device="cuda", generator=g).bfloat16()[:, :K]
    b = torch.randn(N, Kp, device="cuda", generator=g).bfloat16()[:, :K]
    bias = torch.randn(N, device="cuda", generator=g)
    c = gemm_bf16(a, b, bias=bias, relu=True)
    torch.cuda.synchronize()
    ref = _ref(a, b, bias, True)
    torch.testing.assert_close(c, ref, rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("a_mn,b_mn", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 256, 192), (1248, 256, 2048), (257, 128, 1000)])
def test_gemm_mn_major(a_mn, b_mn, M, N, K):
    from deepfm_b200.ops.gemm import gemm_bf16
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    Mp, Np = (M + 7) // 8 * 8, (N + 7) // 8 * 8
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = torch.randn(N, K, device="cuda", generator=g).bfloat16()
    a_in = a
    b_in = b
    if a_mn:
        buf = torch.zeros(K, Mp, device="cuda", dtype=torch.bfloat16)
        buf[:, :M] = a.t()
        a_in = buf[:, :M]
    if b_mn:
        buf = torch.zeros(K, Np, device="cuda", dtype=torch.bfloat16)
        buf[:, :N] = b.t()
        b_in = buf[:, :N]
    c = gemm_bf16(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn)
    torch.cuda.synchronize()
    ref = _ref(a, b, None, False)
    torch.testing.assert_close(c, ref, rtol=2e-3, atol=2e-2)
