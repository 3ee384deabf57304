"""tcgen05 GEMM against the SIMT FFMA GEMM on identical bf16-rounded operands (both fp32
accumulate): differences are accumulation-order only.  Shapes are the engine's call sites."""
import pytest
import torch

from tests.helpers import make_engine

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = make_engine(1, precision="bf16", max_streams=2, kv_capacity=16)
    yield e
    e.close()


def ref(a, w):
    return a.to(torch.bfloat16).float() @ w.to(torch.bfloat16).float().t()


@pytest.mark.parametrize("M,N,K", [
    (64, 3072, 1024), (64, 1024, 4096), (64, 4096, 1024), (1, 1024, 1024), (7, 512, 512), (64, 32, 512),
    (200, 10240, 512), (256, 1024, 1024),                      # swapped (skinny) orientation, split-K
    (1024, 1536, 512), (1024, 512, 2048), (300, 64, 128), (4096, 32, 192), (1000, 640, 512),  # row-major
])
def test_plain_shapes(eng, M, N, K):
    torch.manual_seed(M * 7 + N)
    a, w = torch.randn(M, K), torch.randn(N, K) / K ** 0.5
    want = ref(a, w)
    simt = eng.debug_gemm(0, a, w, M=M, K=K, ld=K)
    tc = eng.debug_gemm(1, a, w, M=M, K=K, ld=K)
    assert torch.allclose(simt, want, atol=2e-3, rtol=1e-3)
    err = float((tc - want).abs().max())
    print(f"M={M} N={N} K={K}: tcgen05 max abs err vs fp64-free reference {err:.2e}")
    assert torch.allclose(tc, want, atol=2e-3, rtol=1e-3), err


@pytest.mark.parametrize("n,T,hist,C,taps,N", [
    (5, 16, 6, 512, 7, 512),      # seanet conv0
    (5, 16, 1, 512, 2, 1536),     # convtr0 (K = 2 C_in, N = S C_out)
    (3, 96, 2, 256, 3, 128),      # res0 conv3
    (3, 480, 1, 128, 2, 256),     # convtr2
    (2, 1920, 2, 64, 3, 32),      # res2 conv3
])
def test_overlapping_row_conv_shapes(eng, n, T, hist, C, taps, N):
    torch.manual_seed(T + C)
    rpb = hist + T
    buf = torch.randn(n, rpb, C)
    w = torch.randn(N, taps * C) / (taps * C) ** 0.5
    K = taps * C
    win = buf.unfold(1, taps, 1).permute(0, 1, 3, 2).reshape(n, rpb - taps + 1, K)   # valid windows
    want = ref(win.reshape(-1, K), w).view(n, rpb - taps + 1, N)[:, :T]
    for backend in (0, 1):
        out = eng.debug_gemm(backend, buf, w, M=n * rpb, K=K, ld=C, rows_per_batch=rpb, valid_rows=T)
        got = out.view(n, rpb, N)[:, :T]
        err = float((got - want).abs().max())
        print(f"conv n={n} T={T} C={C} taps={taps} N={N} backend={backend}: max abs err {err:.2e}")
        assert torch.allclose(got, want, atol=2e-3, rtol=1e-3), (backend, err)
