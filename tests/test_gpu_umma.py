"""Known-answer tests of the tcgen05 conventions (K-panel smem layout + descriptors, TS-form A packing,
cta_group::2 operand split / commit multicast) that the ensemble kernel is built on."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf16_bits(x):
    t = torch.as_tensor(x, dtype=torch.float32).to(torch.bfloat16)
    return t.view(torch.int16).contiguous(), t.to(torch.float32).numpy()


@pytest.mark.parametrize('cg,ts,N,K', [(1, 0, 208, 208), (1, 1, 208, 208), (1, 0, 48, 32), (1, 1, 112, 16),
                                       (2, 0, 208, 208), (2, 1, 208, 208), (2, 1, 112, 208), (2, 1, 96, 208),
                                       (2, 0, 112, 32), (2, 1, 48, 208), (2, 0, 96, 32)])
def test_umma_probe(cg, ts, N, K):
    from mopo_b200 import _lib
    lib = _lib.load_probes()          # the probe build (libmopo_b200_dbg.so, include/mopo_b200_debug.h): not the product library
    M = 512
    rng = np.random.RandomState(N * 7 + K + cg)
    a_bits, a = _bf16_bits(rng.standard_normal((M, K)))
    b_bits, b = _bf16_bits(rng.standard_normal((N, K)))
    dA, dB = a_bits.cuda(), b_bits.cuda()
    D = torch.full((M, N), float('nan'), device='cuda')
    _lib.check(lib.mopo_debug_umma_probe(cg, ts, dA.data_ptr(), dB.data_ptr(), D.data_ptr(), M, N, K,
                                         torch.cuda.current_stream().cuda_stream), lib)
    torch.cuda.synchronize()
    ref = a.astype(np.float64) @ b.astype(np.float64).T
    np.testing.assert_allclose(D.cpu().numpy(), ref, rtol=1e-4, atol=1e-3)
