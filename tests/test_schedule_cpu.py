"""CPU (host code of libmm25 only, no device work): the static schedule of the grouped GEMM
(gemm_tc.cu `gemm_schedule` + the `job_at` walk, restated here from the kernel's formulas) must
hand every block of both problems to exactly one CTA group, and level the per-group cost."""
import ctypes

import pytest


def _schedule(L, mb0, nk0, mb1, nk1, ngroups):
    out = (ctypes.c_int * 5)()
    rc = L.mm25_debug_gemm_schedule(mb0, nk0, mb1, nk1, ngroups, out)
    assert rc == 0
    return list(out)


def _walk(mb0, mb1, ngroups, r, nbA, nbB, nsA, nsB):
    """job_at of gemm_tc.cu / gemm_tc2.cu for every group: lists of (problem, block)."""
    jobs = []
    for grp in range(ngroups):
        A = grp < r
        nb, ns = (nbA, nsA) if A else (nbB, nsB)
        base = grp * nsA if A else r * nsA + (grp - r) * nsB
        mine = []
        k = 0
        while True:
            if k < nb:
                prob, job, lim = 0, grp + k * ngroups, mb0
            else:
                ks = k - nb
                if ks >= ns:
                    break
                prob, job, lim = 1, base + ks, mb1
            if job >= lim:
                break
            mine.append((prob, job))
            k += 1
        jobs.append(mine)
    return jobs


# (mb0, nk0, mb1, nk1, ngroups): cfg2 panel on 148 SMs (1-CTA, d=512: 74 groups of 2), the same
# on CTA pairs (37 groups of 2 pairs, 256-row super-blocks), rank-local shards at P = 2, 8,
# d = 1024 (groups of 4), single problem, fewer blocks than groups, ragged remainders
CASES = [(128, 512, 256, 256, 74), (64, 512, 128, 256, 37), (64, 512, 256, 128, 74),
         (32, 512, 256, 64, 74), (256, 256, 128, 512, 37), (128, 1024, 512, 256, 37),
         (7, 64, 5, 32, 3), (1, 8, 0, 0, 1), (5, 16, 0, 0, 74), (3, 512, 300, 2, 74),
         (100, 3, 1, 700, 9), (74, 10, 74, 10, 74), (75, 10, 73, 10, 74)]


@pytest.mark.parametrize("mb0,nk0,mb1,nk1,ngroups", CASES)
def test_grouped_gemm_schedule_covers_and_balances(lib_built, mb0, nk0, mb1, nk1, ngroups):
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    ngroups = max(1, min(ngroups, mb0 + mb1))  # gemm_grouped clamps the same way
    r, nbA, nbB, nsA, nsB = _schedule(L, mb0, nk0, mb1, nk1, ngroups)
    jobs = _walk(mb0, mb1, ngroups, r, nbA, nbB, nsA, nsB)
    seen = sorted(j for mine in jobs for j in mine)
    assert seen == sorted([(0, b) for b in range(mb0)] + [(1, b) for b in range(mb1)])
    cost = [sum(nk0 if p == 0 else nk1 for p, _ in mine) for mine in jobs]
    ideal = (mb0 * nk0 + mb1 * nk1) / ngroups
    # the makespan is within one block of the ideal level (the unit of work is a whole block)
    assert max(cost) <= ideal + max(nk0, nk1 if mb1 else 0) + 1e-9


def test_grouped_gemm_schedule_cfg2_efficiency(lib_built):
    """the headline shape: 1024 cost units over 74 groups -> 98.8 %; the
    CTA-pair variant keeps the same level."""
    from multimodal_2025_b200 import _lib
    L = _lib.lib()
    for mb0, nk0, mb1, nk1, ng in ((128, 512, 256, 256, 74), (64, 512, 128, 256, 37)):
        r, nbA, nbB, nsA, nsB = _schedule(L, mb0, nk0, mb1, nk1, ng)
        jobs = _walk(mb0, mb1, ng, r, nbA, nbB, nsA, nsB)
        cost = [sum(nk0 if p == 0 else nk1 for p, _ in mine) for mine in jobs]
        eff = (mb0 * nk0 + mb1 * nk1) / ng / max(cost)
        assert eff > 0.985, eff
