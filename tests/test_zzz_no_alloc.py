"""GPU: the allocation-free entry point `mm25_rownorm_colsum_noalloc` (include/mm25.h; no
counterpart in the reference) against the existing
`mm25_rownorm_fwd` through the C-ABI, on two calls in a row that share the library's static
scratch array.  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu

F32, BF16 = 0, 1


def _both(L, x, out_dtype, normalize):
    """(xhat, inv_norm, colsum[B, d]) from the existing entry point and from the new one."""
    B, M, d = x.shape
    st = torch.cuda.current_stream().cuda_stream
    dt = BF16 if x.dtype == torch.bfloat16 else F32
    odt = torch.bfloat16 if out_dtype == BF16 else torch.float32
    outs = []
    for new in (False, True):
        xhat = torch.empty(B, M, d, dtype=odt, device="cuda")
        inv = torch.empty(B, M, dtype=torch.float32, device="cuda")
        if new:
            cs = torch.full((B, d), float("nan"), device="cuda")
            rc = L.mm25_rownorm_colsum_noalloc(x.data_ptr(), dt, xhat.data_ptr(), out_dtype,
                                               inv.data_ptr(), cs.data_ptr(), B, M, d,
                                               int(normalize), 1e-12, st)
        else:
            cs = torch.empty(int(L.mm25_rownorm_colsum_floats(B, M, d)), device="cuda")
            rc = L.mm25_rownorm_fwd(x.data_ptr(), dt, xhat.data_ptr(), out_dtype, inv.data_ptr(),
                                    cs.data_ptr(), B, M, d, int(normalize), 1e-12, st)
            cs = cs[:B * d].reshape(B, d)
        assert rc == 0, rc
        outs.append((xhat, inv, cs))
    torch.cuda.synchronize()
    return outs


def test_noalloc_rownorm_colsum_matches_existing_entry_twice(lib_built):
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(0)
    # two calls in a row; the second is smaller, so the static array still holds partials of the
    # first beyond what the second overwrites
    cases = [(torch.randn(2, 4096, 512, generator=g), BF16, True),
             (torch.randn(1, 320, 256, generator=g).bfloat16(), BF16, True),
             (torch.randn(1, 1000, 1024, generator=g) * 3.0, F32, False)]
    for x, out_dtype, normalize in cases:
        (xh0, inv0, cs0), (xh1, inv1, cs1) = _both(L, x.cuda(), out_dtype, normalize)
        assert torch.equal(xh0, xh1) and torch.equal(inv0, inv1)
        assert torch.isfinite(cs1).all()
        # same kernels, same fixed summation order: fp32 tolerance of the suite (1e-5) is ample
        assert (cs1 - cs0).abs().max().item() <= 1e-5 * max(1.0, cs0.abs().max().item())
        ref = xh0.float().sum(dim=1)
        assert (cs1 - ref).abs().max().item() <= 1e-3 * max(1.0, ref.abs().max().item())


def test_noalloc_rejects_sizes_beyond_the_static_scratch(lib_built):
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    x = torch.zeros(1, 64, 128, device="cuda")
    o = torch.zeros(1, 64, 128, device="cuda")
    cs = torch.zeros(1, 128, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    # B * d * min(ceil(M / 8), 740) = 8 * 2048 * 740 floats > MM25_NOALLOC_MAX_FLOATS
    assert L.mm25_rownorm_colsum_noalloc_floats(8, 32768, 2048) > 740 * 1024
    rc = L.mm25_rownorm_colsum_noalloc(x.data_ptr(), F32, o.data_ptr(), F32, None, cs.data_ptr(),
                                       8, 32768, 2048, 1, 1e-12, st)
    assert rc == -4                                   # MM25_ERR_WORKSPACE, nothing launched
    rc = L.mm25_rownorm_colsum_noalloc(x.data_ptr(), F32, o.data_ptr(), F32, None, cs.data_ptr(),
                                       1, 64, 100, 1, 1e-12, st)
    assert rc == -1                                   # MM25_ERR_SHAPE: d % 128 != 0
    rc = L.mm25_rownorm_colsum_noalloc(x.data_ptr(), F32, o.data_ptr(), F32, None, None,
                                       1, 64, 128, 1, 1e-12, st)
    assert rc == -7                                   # MM25_ERR_ARG: colsum is an output
    rc = L.mm25_rownorm_colsum_noalloc(x.data_ptr(), F32, o.data_ptr(), F32, None, cs.data_ptr(),
                                       1, 64, 128, 1, 1e-12, st)
    assert rc == 0
    torch.cuda.synchronize()
