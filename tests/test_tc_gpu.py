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


@pytest.mark.parametrize("B,M,N,d,off,ls", [
    (1, 256, 256, 256, 0, 2.6592),
    (1, 384, 1024, 512, 256, 2.6592),    # rank-local rectangular block, shifted positives
    (3, 256, 512, 256, 0, 3.2),          # batched (vCLIP samples)
    (1, 2176, 2304, 512, 0, 2.0),        # not a multiple of the 2048-column block
    (1, 256, 256, 256, 0, 4.6052),       # spread too large for the static shift: flag stays 0
    # shapes of the fused-transform GEMM (csrc/gemm_te.cu: M, N, d multiples of 256)
    (1, 512, 1024, 512, 256, 2.6592),    # rectangular, shifted positives
    (2, 512, 512, 1024, 0, 3.0),         # d = 1024: two 512-column tiles; batched
    (1, 768, 1024, 256, 0, 2.0),         # d = 256: one MMA per k-step, two accumulator stages
    (1, 768, 1280, 768, 0, 2.6592),      # d = 768: three 256-column tiles
    (1, 4096, 4096, 512, 0, 2.6592),     # many pairs, tiles split between two pairs (red.add)
    (1, 512, 512, 512, 0, 4.6052),       # flag 0 on an eligible shape: recompute + copy mode
])
def test_tc_backward_kept_e_matches_recompute(lib_built, B, M, N, d, off, ls):
    """the kept-E backward (forward stores E, elementwise E -> G', dscale from a . dA) against
    the S-tile recompute backward on the same inputs, and E itself against its definition."""
    from multimodal_2025_b200 import ops
    a, b = make(B, M, N, d, 7)
    a = torch.nn.functional.normalize(a.float(), dim=-1).bfloat16()
    b = torch.nn.functional.normalize(b.float(), dim=-1).bfloat16()
    scale = torch.tensor([ls], device="cuda")
    bound = ops.DOT_BOUND_NORMALIZED
    row_lse, cmax, csum, diag, e_keep, e_flag = ops.sim_fwd_keep(
        a, b, scale, ops.SCALE_LOG, 4.6052, off, True, ops.IMPL_TC, bound)
    plain = ops.sim_fwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, off,
                        True, ops.IMPL_TC, bound)
    for got, ref in zip((row_lse, cmax, csum, diag), plain):
        assert torch.equal(got, ref)  # keeping E does not change the statistics
    s = float(torch.tensor(ls).clamp(max=4.6052).exp())
    fast = 2 * bound * s * 1.4426950408889634 <= 96.0
    assert int(e_flag.item()) == (1 if fast else 0)
    if fast:
        z = s * (a.float() @ b.float().transpose(1, 2))
        e_ref = torch.exp(z - bound * s).reshape(B * M, N)
        rel = ((e_keep.float() - e_ref).abs() / e_ref.clamp_min(1e-30)).max().item()
        assert rel < 1.5e-2, rel  # bf16 rounding (2^-8) + ex2.approx
    col_lse = (cmax + csum.log())
    nn_ = float(N)
    w = (0.5 / nn_, 0.5 / nn_, 0.9 / nn_, 0.1 / (nn_ * nn_))
    gout = torch.linspace(0.5, 1.5, B, device="cuda")
    ref = ops.sim_bwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, off,
                      row_lse, col_lse, *w, gout, True, True, False, ops.IMPL_TC, bound)
    got = ops.sim_bwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, off,
                      row_lse, col_lse, *w, gout, True, True, False, ops.IMPL_TC, bound,
                      e_keep, e_flag)
    # fp32 closed form on the same bf16 operands: the bf16 contract (2e-2) holds
    # for both paths; E's second bf16 rounding makes the kept-E path the looser of the two
    af, bf = a.float(), b.float()
    z = s * (af @ bf.transpose(1, 2))
    r = torch.logsumexp(z, dim=2, keepdim=True)
    c = torch.logsumexp(z, dim=1, keepdim=True)
    G = w[0] * (z - r).exp() + w[1] * (z - c).exp() - w[3]
    idx = torch.arange(M, device="cuda")
    G[:, idx, idx + off] -= w[2]
    G = G * gout.reshape(B, 1, 1) * s
    exact = (G @ bf, G.transpose(1, 2) @ af)
    for k in (0, 1):
        den = exact[k].norm().item()
        e_rec = (ref[k] - exact[k]).norm().item() / den
        e_kept = (got[k] - exact[k]).norm().item() / den
        assert e_rec <= 2e-2 and e_kept <= 2e-2, (k, e_rec, e_kept)
    assert abs(got[2].item() - ref[2].item()) <= 2e-2 * abs(ref[2].item()) + 1e-7
    # only da requested: dscale still comes out (from a . dA); only db: from b . dB
    e2 = ops.sim_fwd_keep(a, b, scale, ops.SCALE_LOG, 4.6052, off, True, ops.IMPL_TC, bound)
    only_db = ops.sim_bwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN,
                          off, row_lse, col_lse, *w, gout, False, True, False, ops.IMPL_TC, bound,
                          e2[4], e2[5])
    assert abs(only_db[2].item() - ref[2].item()) <= 2e-2 * abs(ref[2].item()) + 1e-7
    assert (only_db[1] - exact[1]).norm().item() <= 2e-2 * exact[1].norm().item()


@pytest.mark.parametrize("Me,Ne,Ng,batch,off,flag", [
    (256, 256, 256, 1, 0, 1), (512, 768, 512, 1, 0, 1), (256, 1024, 512, 1, 512, 1),
    (768, 512, 1024, 1, -256, 1), (512, 512, 512, 3, 0, 1), (1024, 2048, 512, 1, 0, 0),
    (2048, 2048, 512, 1, 0, 1), (512, 256, 768, 2, 0, 1),
])
def test_gemm_te_transform_on_load_vs_torch(lib_built, Me, Ne, Ng, batch, off, flag):
    """C-ABI test entry of csrc/gemm_te.cu: C0 = T(E) B0, C1 = T(E)^T B1 with
    T(E) = E (rowf + colf) - k_all, T(E) = gdiag on the positives, against the same arithmetic in
    torch (bf16-rounded T(E), fp32 products)."""
    import ctypes
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(Me + Ne + Ng + batch)
    E = torch.rand(batch * Me, Ne, generator=g).cuda().bfloat16()
    rowf = (torch.rand(batch * Me, generator=g) * 2e-3).cuda()
    colf = (torch.rand(batch * Ne, generator=g) * 2e-3).cuda()
    consts = torch.tensor([[3e-5 * (z + 1), 0.0] for z in range(batch)]).cuda().contiguous()
    gdiag = (torch.rand(batch * Me, generator=g) * 1e-3 - 5e-4).cuda()   # exact G' of positives
    e_flag = torch.tensor([flag], dtype=torch.int32, device="cuda")
    B0 = torch.randn(batch * Ne, Ng, generator=g).cuda().bfloat16()
    B1 = torch.randn(batch * Me, Ng, generator=g).cuda().bfloat16()
    C0 = torch.full((batch * Me, Ng), 7.0, device="cuda")   # the entry zeroes C itself
    C1 = torch.full((batch * Ne, Ng), 7.0, device="cuda")
    rc = L.mm25_debug_gemm_te(E.data_ptr(), Me, Ne, rowf.data_ptr(), colf.data_ptr(),
                              consts.data_ptr(), gdiag.data_ptr(), e_flag.data_ptr(), off, B0.data_ptr(),
                              C0.data_ptr(), B1.data_ptr(), C1.data_ptr(), Ng, batch,
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    for z in range(batch):
        Ez = E[z * Me:(z + 1) * Me].float()
        if flag:
            T = Ez * (rowf[z * Me:(z + 1) * Me, None] + colf[None, z * Ne:(z + 1) * Ne]) \
                - consts[z, 0]
            r = torch.arange(Me, device="cuda")
            ok = (r + off >= 0) & (r + off < Ne)
            T[r[ok], (r + off)[ok]] = gdiag[z * Me:(z + 1) * Me][ok]
            T = T.bfloat16().float()
        else:
            T = Ez
        r0 = T @ B0[z * Ne:(z + 1) * Ne].float()
        r1 = T.t() @ B1[z * Me:(z + 1) * Me].float()
        for got, ref in ((C0[z * Me:(z + 1) * Me], r0), (C1[z * Ne:(z + 1) * Ne], r1)):
            # identical bf16 operands up to the last-ulp rounding of T: fp32 accumulation error
            assert (got - ref).norm().item() <= 3e-3 * ref.norm().item(), (z, flag)
    # one-sided calls (only da / only db requested)
    C0b = torch.empty_like(C0)
    rc = L.mm25_debug_gemm_te(E.data_ptr(), Me, Ne, rowf.data_ptr(), colf.data_ptr(),
                              consts.data_ptr(), gdiag.data_ptr(), e_flag.data_ptr(), off, B0.data_ptr(),
                              C0b.data_ptr(), None, None, Ng, batch,
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    C1b = torch.empty_like(C1)
    rc = L.mm25_debug_gemm_te(E.data_ptr(), Me, Ne, rowf.data_ptr(), colf.data_ptr(),
                              consts.data_ptr(), gdiag.data_ptr(), e_flag.data_ptr(), off, None, None,
                              B1.data_ptr(), C1b.data_ptr(), Ng, batch,
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    # same products; the stream-K split points differ with the amount of work in the launch, so
    # split tiles may differ in the last fp32 bits
    assert (C0b - C0).norm().item() <= 1e-5 * C0.norm().item()
    assert (C1b - C1).norm().item() <= 1e-5 * C1.norm().item()
    # the same launch twice: tiles split between more than two pairs (short K ranges on small
    # problems) are added in arrival order, so repeat runs agree to fp32 rounding, not bitwise
    C1c = torch.empty_like(C1)
    rc = L.mm25_debug_gemm_te(E.data_ptr(), Me, Ne, rowf.data_ptr(), colf.data_ptr(),
                              consts.data_ptr(), gdiag.data_ptr(), e_flag.data_ptr(), off, None, None,
                              B1.data_ptr(), C1c.data_ptr(), Ng, batch,
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert (C1c - C1b).norm().item() <= 1e-6 * C1b.norm().item()


def test_gemm_te_is_bit_reproducible_when_tiles_split_in_two(lib_built):
    """every job's K range (nk = 4 k-blocks) is shorter than a pair's share (8), so a tile is the
    sum of at most two partials added onto zero: 0 + a + b is order-independent."""
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(3)
    Me, Ne, Ng = 256, 19456, 256
    E = torch.rand(Me, Ne, generator=g).cuda().bfloat16()
    rowf = (torch.rand(Me, generator=g) * 2e-3).cuda()
    colf = (torch.rand(Ne, generator=g) * 2e-3).cuda()
    consts = torch.tensor([1e-5, 0.0]).cuda()
    one = torch.ones(1, dtype=torch.int32, device="cuda")
    B1 = torch.randn(Me, Ng, generator=g).cuda().bfloat16()
    outs = []
    for _ in range(3):
        C1 = torch.empty(Ne, Ng, device="cuda")
        rc = L.mm25_debug_gemm_te(E.data_ptr(), Me, Ne, rowf.data_ptr(), colf.data_ptr(),
                                  consts.data_ptr(), None, one.data_ptr(), 0, None, None,
                                  B1.data_ptr(), C1.data_ptr(), Ng, 1,
                                  torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        outs.append(C1)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    T = (E.float() * (rowf[:, None] + colf[None, :]) - consts[0]).bfloat16().float()
    ref = T.t() @ B1.float()
    assert (outs[0] - ref).norm().item() <= 3e-3 * ref.norm().item()


def test_gemm_te_rejects_bad_shapes(lib_built):
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    E = torch.zeros(256, 256, device="cuda", dtype=torch.bfloat16)
    f = torch.zeros(256, device="cuda")
    B0 = torch.zeros(256, 256, device="cuda", dtype=torch.bfloat16)
    C0 = torch.zeros(256, 256, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    # Me not a multiple of 256
    rc = L.mm25_debug_gemm_te(E.data_ptr(), 128, 256, f.data_ptr(), f.data_ptr(), None, None, None,
                              0, B0.data_ptr(), C0.data_ptr(), None, None, 256, 1, st)
    assert rc != 0
    rc = L.mm25_debug_gemm_te(None, 256, 256, f.data_ptr(), f.data_ptr(), None, None, None, 0,
                              B0.data_ptr(), C0.data_ptr(), None, None, 256, 1, st)
    assert rc != 0
