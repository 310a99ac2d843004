"""tcgen05 path: descriptor probe first, then the fused bf16 chain against the fp32 CUDA path and the oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda", 0)


def _probe(dev, mode, N, K, seed=0):
    from codenerf_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(seed)
    if mode == 0:
        A, B = torch.randn(128, K, generator=g), torch.randn(N, K, generator=g)
        want = A.bfloat16().double() @ B.bfloat16().double().t()
    else:
        A, B = torch.randn(K, 128, generator=g), torch.randn(K, N, generator=g)
        want = A.bfloat16().double().t() @ B.bfloat16().double()
    Ad, Bd = A.to(dev), B.to(dev)
    D = torch.full((128, N), float("nan"), device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    p = lambda t: C.c_void_p(t.data_ptr())
    _lib.check(lib.cnrf_debug_umma(mode, p(Ad), p(Bd), p(D), N, K, p(err), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
               "cnrf_debug_umma")
    torch.cuda.synchronize()
    assert int(err.item()) == 0, "mbarrier wait timed out (MMA never committed)"
    return D.cpu().double(), want


@pytest.mark.parametrize("N,K", [(256, 64), (256, 256), (64, 128), (16, 64)])
def test_umma_k_major(dev, N, K):
    got, want = _probe(dev, 0, N, K)
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("N,K", [(256, 64), (256, 128), (64, 64), (128, 192)])
def test_umma_mn_major(dev, N, K):
    got, want = _probe(dev, 1, N, K)
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=1e-4, atol=1e-3)
