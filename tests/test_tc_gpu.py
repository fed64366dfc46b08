"""GPU: the tcgen05 / TMEM / TMA forward kernel against an fp32 torch evaluation of the same
bf16 operands (statistics level: row LSE, column LSE, positive logits)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def ref_stats(a, b, s, diag_offset):
    z = s * (a.float() @ b.float().transpose(-1, -2))
    r = torch.logsumexp(z, dim=-1)
    c = torch.logsumexp(z, dim=-2)
    M = a.shape[-2]
    idx = torch.arange(M, device=a.device) + diag_offset
    dg = z[..., torch.arange(M, device=a.device), idx]
    return r, c, dg


def make(B, M, N, d, seed, spread=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    a = torch.nn.functional.normalize(torch.randn(B, M, d, generator=g), dim=-1)
    b = torch.nn.functional.normalize(torch.randn(B, N, d, generator=g), dim=-1)
    k = min(M, N)
    b[:, :k] = torch.nn.functional.normalize(a[:, :k] * spread + b[:, :k], dim=-1)
    return a.cuda().bfloat16(), b.cuda().bfloat16()


@pytest.mark.parametrize("B,M,N,d,off,ls,force_slow", [
    (1, 128, 256, 64, 0, 2.0, 0),
    (1, 256, 256, 512, 0, 2.6592, 0),
    (1, 1024, 1024, 512, 0, 4.6052, 0),
    (2, 256, 512, 128, 128, 3.0, 0),
    (1, 512, 2048, 256, 1024, 4.6052, 0),
    (1, 256, 256, 512, 0, 4.6052, 1),
    (3, 128, 256, 64, 0, 1.0, 1),
    (1, 4096, 4096, 512, 0, 4.0, 0),
])
def test_tc_forward_stats(lib_built, B, M, N, d, off, ls, force_slow):
    from multimodal_2025_b200 import ops, _lib
    h = ctypes.CDLL(_lib.LIB_PATH)
    h.mm25_debug_force_slow_epilogue(force_slow)
    try:
        a, b = make(B, M, N, d, 3, spread=2.0)
        scale = torch.tensor([ls], device="cuda")
        row_lse, cmax, csum, diag = ops.sim_fwd(a, b, None, None, None, scale, ops.SCALE_LOG,
                                                4.6052, ops.SIM_PLAIN, off, True, ops.IMPL_TC)
        torch.cuda.synchronize()
    finally:
        h.mm25_debug_force_slow_epilogue(0)
    s = float(torch.tensor(min(ls, 4.6052)).exp())
    r, c, dg = ref_stats(a, b, s, off)
    col_lse = cmax + csum.log()
    assert torch.allclose(diag, dg, atol=2e-3 * s, rtol=0), (diag - dg).abs().max()
    assert (row_lse - r).abs().max() < 2e-3
    assert (col_lse - c).abs().max() < 2e-3


@pytest.mark.parametrize("B,M,N,d,off,ls", [
    (1, 256, 256, 512, 0, 2.6592),
    (2, 256, 512, 128, 128, 3.4),
    (1, 1024, 2048, 256, 512, 1.0),
    (1, 256, 256, 64, 0, 4.6052),   # bound given but s too large: must take the general path
])
def test_tc_forward_stats_static_shift(lib_built, B, M, N, d, off, ls):
    """caller-supplied |a.b| bound -> static-shift epilogue (one exp, no running max)."""
    from multimodal_2025_b200 import ops
    a, b = make(B, M, N, d, 4, spread=2.0)
    scale = torch.tensor([ls], device="cuda")
    row_lse, cmax, csum, diag = ops.sim_fwd(a, b, None, None, None, scale, ops.SCALE_LOG,
                                            4.6052, ops.SIM_PLAIN, off, True, ops.IMPL_TC,
                                            ops.DOT_BOUND_NORMALIZED)
    s = float(torch.tensor(min(ls, 4.6052)).exp())
    r, c, dg = ref_stats(a, b, s, off)
    col_lse = cmax + csum.log()
    assert torch.allclose(diag, dg, atol=2e-3 * s, rtol=0)
    assert (row_lse - r).abs().max() < 2e-3
    assert (col_lse - c).abs().max() < 2e-3


def test_tc_matches_simt_on_adversarial_range(lib_built):
    """duplicates (cos = 1) next to unrelated pairs at s = 100: > 87 nats inside one block."""
    from multimodal_2025_b200 import ops
    a, b = make(1, 256, 256, 128, 5, spread=0.0)
    b[:, ::2] = a[:, ::2]
    b[:, 1::4] = -a[:, 1::4]
    scale = torch.tensor([4.6052], device="cuda")
    out_tc = ops.sim_fwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, 0,
                         True, ops.IMPL_TC)
    out_si = ops.sim_fwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, 0,
                         True, ops.IMPL_SIMT)
    assert (out_tc[0] - out_si[0]).abs().max() < 2e-3
    c_tc = out_tc[1] + out_tc[2].log()
    c_si = out_si[1] + out_si[2].log()
    assert torch.isfinite(c_tc).all()
    assert (c_tc - c_si).abs().max() < 2e-3


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("Mg,Ng,K,acc", [(128, 256, 64, 0), (256, 512, 1024, 0), (512, 256, 320, 1)])
def test_tc_gemm_layouts(lib_built, a_mn, b_mn, Mg, Ng, K, acc):
    """the bf16 tcgen05 GEMM of the backward, all four operand-major combinations."""
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cpu").manual_seed(Mg + K)
    A = torch.randn(Mg, K, generator=g).cuda().bfloat16()
    B = torch.randn(Ng, K, generator=g).cuda().bfloat16()
    ref = 0.5 * (A.float() @ B.float().t())
    C = torch.full((Mg, Ng), 1.0, device="cuda")
    if acc:
        ref = ref + 1.0
    Am = A.t().contiguous() if a_mn else A.contiguous()
    Bm = B.t().contiguous() if b_mn else B.contiguous()
    rc = L.mm25_debug_gemm_bf16(Am.data_ptr(), a_mn, Bm.data_ptr(), b_mn, C.data_ptr(), Mg, Ng,
                                K, 0.5, acc, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    err = (C - ref).abs().max().item()
    assert err < 1e-2 * max(1.0, ref.abs().max().item() / 10), err


@pytest.mark.parametrize("M0,K0,M1,K1,Ng", [(256, 2048, 2048, 256, 512), (128, 512, 512, 128, 256),
                                            (1024, 4096, 4096, 1024, 512), (384, 640, 640, 384, 1024),
                                            (512, 16384, 16384, 512, 512)])  # last: split-K
def test_tc_gemm_grouped_pair(lib_built, M0, K0, M1, K1, Ng):
    """dA = G . Y (K-major x MN-major) and dB = G^T . X (MN x MN) in one scheduled launch."""
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cpu").manual_seed(M0 + K0)
    G = torch.randn(M0, K0, generator=g).cuda().bfloat16()      # [R, N]
    Y = torch.randn(K0, Ng, generator=g).cuda().bfloat16()      # [N, d]
    X = torch.randn(K1, Ng, generator=g).cuda().bfloat16()      # [R, d]
    assert M1 == K0 and K1 == M0
    C0 = torch.zeros(M0, Ng, device="cuda")
    C1 = torch.zeros(M1, Ng, device="cuda")
    rc = L.mm25_debug_gemm_grouped(G.data_ptr(), 0, Y.data_ptr(), 1, C0.data_ptr(), M0, K0,
                                   G.data_ptr(), 1, X.data_ptr(), 1, C1.data_ptr(), M1, K1, Ng,
                                   torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    r0 = G.float() @ Y.float()
    r1 = G.float().t() @ X.float()
    assert (C0 - r0).abs().max().item() < 1e-2 * max(1.0, r0.abs().max().item() / 10)
    assert (C1 - r1).abs().max().item() < 1e-2 * max(1.0, r1.abs().max().item() / 10)
