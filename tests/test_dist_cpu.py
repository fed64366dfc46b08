"""CPU, world_size 2 over gloo: the sharded-loss collective logic (all-gather, column-partial
merge, reduce-scatter, scalar all-reduces) against the single-process oracle."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _np(obj):
    """tensors -> numpy (recursively): results cross the process boundary BY VALUE.  A tensor put
    on a multiprocessing queue travels as a shared-memory handle that dies with the sender, so a
    worker that exits before the parent has rebuilt it raises FileNotFoundError in the parent."""
    if torch.is_tensor(obj):
        return ("__tensor__", obj.detach().cpu().numpy())
    if isinstance(obj, dict):
        return {k: _np(v) for k, v in obj.items()}
    if isinstance(obj, (tuple, list)):
        return type(obj)(_np(v) for v in obj)
    return obj


def _th(obj):
    if isinstance(obj, tuple) and len(obj) == 2 and isinstance(obj[0], str) and \
            obj[0] == "__tensor__":
        return torch.from_numpy(obj[1].copy())
    if isinstance(obj, dict):
        return {k: _th(v) for k, v in obj.items()}
    if isinstance(obj, (tuple, list)):
        return type(obj)(_th(v) for v in obj)
    return obj


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run_world(target, world, args, attempts=2):
    """Spawn `world` gloo ranks and collect one result per rank; a failed rendezvous (port taken
    between the probe and the bind, slow start-up on a loaded box) is retried on a fresh port."""
    import queue as _queue
    last = None
    for _ in range(attempts):
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=target, args=(r, world, port, q) + tuple(args))
                 for r in range(world)]
        for p in procs:
            p.start()
        try:
            res = sorted([_th(q.get(timeout=180)) for _ in range(world)], key=lambda t: t[0])
            ok = True
        except _queue.Empty as e:
            res, ok, last = None, False, e
        for p in procs:
            p.join(timeout=60)
            if p.is_alive():
                p.terminate()
            ok = ok and p.exitcode == 0
        if ok:
            return res
    raise AssertionError("gloo world failed after %d attempts: %r" % (attempts, last))


def _worker(rank, world, port, q, eps, n, d, ls):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cpu_backend
        from multimodal_2025_b200.dist import ShardedClipLoss
        from oracle import losses_oracle as O
        x, y = O.synth_embeddings(n * world, d, 7, correlated=True)
        x = x * 1.7  # exercise the normalise backward
        sl = slice(rank * n, (rank + 1) * n)
        xl = x[sl].clone().requires_grad_(True)
        yl = y[sl].clone().requires_grad_(True)
        lsp = torch.tensor(ls, requires_grad=True)
        crit = ShardedClipLoss(eps, backend=cpu_backend)
        loss = crit(xl, yl, lsp)
        (2.0 * loss).backward()
        q.put(_np((rank, float(loss.detach()), xl.grad.clone(), yl.grad.clone(), float(lsp.grad))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("eps,ls", [(0.1, 2.6592), (0.0, 5.0)])
def test_sharded_clip_world2_gloo(eps, ls):
    from oracle import losses_oracle as O
    world, n, d = 2, 48, 32
    res = _run_world(_worker, world, (eps, n, d, ls))
    x, y = O.synth_embeddings(n * world, d, 7, correlated=True)
    x = x * 1.7
    ref, (rx, ry, rl) = O.loss_and_grads(lambda a, b, c: 2.0 * O.clip_loss(a, b, c, eps, True),
                                         [x, y, torch.tensor(ls)])
    for rank, loss, gx, gy, gl in res:
        assert abs(2.0 * loss - float(ref)) < 1e-5 * abs(float(ref))
        sl = slice(rank * n, (rank + 1) * n)
        assert (gx - rx[sl]).norm() < 1e-4 * rx[sl].norm()
        assert (gy - ry[sl]).norm() < 1e-4 * ry[sl].norm()
        assert abs(gl - float(rl)) <= 1e-4 * max(abs(float(rl)), 1e-6)


def _worker_align(rank, world, port, q, semantic, n, d, dsem):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cpu_backend
        from multimodal_2025_b200.dist import ShardedAlignClipLoss, ShardedClipInModalityLoss
        from oracle import losses_oracle as O
        x, y = O.synth_embeddings(n * world, d, 11, correlated=True)
        sem = torch.randn(n * world, dsem, generator=torch.Generator().manual_seed(4))
        sl = slice(rank * n, (rank + 1) * n)
        xl = x[sl].clone().requires_grad_(True)
        yl = y[sl].clone().requires_grad_(True)
        sc = torch.tensor(14.285, requires_grad=True)
        crit = ShardedClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=semantic,
                                         separate_image=semantic, backend=cpu_backend)
        out = crit(xl, yl, sc, output_dict=True, semantic_features=sem[sl] if semantic else None)
        out["total_loss"].backward()
        plain = ShardedAlignClipLoss(backend=cpu_backend)(xl.detach(), yl.detach(), sc.detach())
        q.put(_np((rank, float(out["total_loss"].detach()), float(out["clip_loss"].detach()),
               float(out["inModality_loss"].detach()), xl.grad.clone(), yl.grad.clone(),
               float(sc.grad), float(plain))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("semantic", [True, False])
def test_sharded_alignclip_world2_gloo(semantic):
    """ShardedClipInModalityLoss / ShardedAlignClipLoss on 2 ranks against the single-process
    oracle on the concatenated batch (reference loss.py:221-273 restated in oracle/)."""
    from oracle import losses_oracle as O
    world, n, d, dsem = 2, 40, 24, 12
    res = _run_world(_worker_align, world, (semantic, n, d, dsem))
    x, y = O.synth_embeddings(n * world, d, 11, correlated=True)
    sem = torch.randn(n * world, dsem, generator=torch.Generator().manual_seed(4))
    s = torch.tensor(14.285)
    fn = lambda a, b, c: O.inmodality_loss(a, b, c, sem if semantic else None,  # noqa: E731
                                           nl_semantic_supervision=semantic,
                                           separate_image=semantic)
    tot, clip, inm = fn(x, y, s)
    ref, (rx, ry, rs) = O.loss_and_grads(lambda a, b, c: fn(a, b, c)[0], [x, y, s])
    oc = O.openclip_loss(x, y, s)
    for rank, ltot, lclip, linm, gx, gy, gs, lplain in res:
        assert abs(ltot - float(tot)) < 1e-5 * abs(float(tot))
        assert abs(lclip - float(clip)) < 1e-5 * abs(float(clip))
        # CE(s I I^T) ~ 3e-3 is a difference of O(10) fp32 terms: absolute floor 1e-6
        assert abs(linm - float(inm)) < 1e-5 * abs(float(inm)) + 1e-6
        assert abs(lplain - float(oc)) < 1e-5 * abs(float(oc))
        sl = slice(rank * n, (rank + 1) * n)
        assert (gx - rx[sl]).norm() < 1e-4 * rx[sl].norm()
        assert (gy - ry[sl]).norm() < 1e-4 * ry[sl].norm()
        assert abs(gs - float(rs)) <= 1e-4 * max(abs(float(rs)), 1e-6)


def _worker_retrieval(rank, world, port, q, seed):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cpu_backend
        from conftest import load_golden
        from multimodal_2025_b200 import dist as mdist
        g = torch.Generator().manual_seed(seed)
        sizes = [37, 23]                       # ragged gallery shards
        off = sum(sizes[:rank])
        scores = torch.randn(9, sum(sizes), generator=g)
        scores[:5] = torch.round(scores[:5] * 2) / 2          # heavy ties in the first rows
        scores[8, 40:] = float("-inf")
        local = scores[:, off:off + sizes[rank]].contiguous()
        out = {}
        for k in (1, 4, 30, 60):               # 30 > the smaller shard, 60 = the whole gallery
            idx, val = mdist.sharded_topk(local, k, backend=cpu_backend)
            out[k] = (idx, val)
        # top_k_score on the reference's golden embeddings, texts sharded 170 / 130
        G = load_golden("topk.npz")[""]
        texts, images = torch.from_numpy(G["texts"]), torch.from_numpy(G["images"])
        cut = [0, 170, 300]
        tl = texts[cut[rank]:cut[rank + 1]]
        owner = torch.arange(300)[cut[rank]:cut[rank + 1]] // 3
        score = {k: mdist.sharded_top_k_score(tl, owner, images, k, backend=cpu_backend)
                 for k in (1, 5, 10, 100)}
        # paired ranks of a row-sharded batch
        n, d = 21, 16
        x = torch.randn(n * world, d, generator=torch.Generator().manual_seed(3))
        y = x + 0.7 * torch.randn(n * world, d, generator=torch.Generator().manual_seed(4))
        sl = slice(rank * n, (rank + 1) * n)
        i2t, t2i = mdist.sharded_paired_ranks(x[sl], y[sl], 7.5, backend=cpu_backend)
        q.put(_np((rank, out, score, i2t, t2i)))
    finally:
        dist.destroy_process_group()


def test_sharded_retrieval_world2_gloo():
    """sharded_topk == exact top-k of the concatenated gallery (stable ties, ragged shards,
    k larger than a shard), sharded_top_k_score == the reference's golden counts, and
    sharded_paired_ranks == the sort-based oracle on the concatenated batch."""
    import cpu_backend
    from conftest import load_golden
    from oracle import metrics_oracle as MO
    world, seed = 2, 17
    res = _run_world(_worker_retrieval, world, (seed,))
    g = torch.Generator().manual_seed(seed)
    scores = torch.randn(9, 60, generator=g)
    scores[:5] = torch.round(scores[:5] * 2) / 2
    scores[8, 40:] = float("-inf")
    G = load_golden("topk.npz")
    n, d = 21, 16
    x = torch.randn(n * world, d, generator=torch.Generator().manual_seed(3))
    y = x + 0.7 * torch.randn(n * world, d, generator=torch.Generator().manual_seed(4))
    s = torch.cat([7.5 * x[r * n:(r + 1) * n] @ y.t() for r in range(world)])
    ar = torch.arange(n * world)
    want_i2t = MO.ranks_by_sort(s, ar)
    want_t2i = MO.ranks_by_sort(s.t().contiguous(), ar)
    for rank, out, score, i2t, t2i in res:
        for k, (idx, val) in out.items():
            widx, wval = cpu_backend.topk_rows(scores, k)
            assert torch.equal(idx, widx), k
            assert torch.equal(val, wval), k
            assert torch.equal(val[5:8], torch.topk(scores[5:8], k, dim=1).values)
        for k, (acc, info) in score.items():
            assert info["correct"] == int(G["k%d" % k]["correct"]) and info["total"] == 100
            assert acc == float(G["k%d" % k]["acc"])
        sl = slice(rank * n, (rank + 1) * n)
        assert torch.equal(i2t.long(), want_i2t[sl])
        assert torch.equal(t2i.long(), want_t2i[sl])


def test_cloob_host_logic_with_standin_vs_oracle():
    """single process: the CLOOB autograd functions (losses/cloobLoss.py) driven through the
    dense stand-in ops reproduce the oracle's loss and gradients — covers the plumbing the
    sharded loss relies on (rectangular retrievals, diag offset, n_norm)."""
    import cpu_backend
    from multimodal_2025_b200.losses.cloobLoss import _HopfieldFn, _InfoLoobFn
    from oracle import losses_oracle as O
    x, y = O.synth_embeddings(40, 24, 5, correlated=True)
    ref, (rx, ry) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0), [x, y])
    xl, yl = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
    H = lambda q, k: _HopfieldFn.apply(q, k, 15.0, cpu_backend)  # noqa: E731
    li = _InfoLoobFn.apply(H(xl, xl), H(yl, xl), 30.0, 0, None, cpu_backend)
    lt = _InfoLoobFn.apply(H(yl, yl), H(xl, yl), 30.0, 0, None, cpu_backend)
    loss = (li + lt) / 2
    loss.backward()
    assert abs(float(loss.detach()) - float(ref)) < 1e-5 * abs(float(ref))
    assert (xl.grad - rx).norm() < 1e-4 * rx.norm()
    assert (yl.grad - ry).norm() < 1e-4 * ry.norm()


def _worker_cloob(rank, world, port, q, n, d):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cpu_backend
        from multimodal_2025_b200.dist import ShardedCLOOBLoss
        from oracle import losses_oracle as O
        x, y = O.synth_embeddings(n * world, d, 9, correlated=True)
        sl = slice(rank * n, (rank + 1) * n)
        xl = x[sl].clone().requires_grad_(True)
        yl = y[sl].clone().requires_grad_(True)
        loss = ShardedCLOOBLoss(30.0, 15.0, "cpu", backend=cpu_backend)(xl, yl)
        (3.0 * loss).backward()
        q.put(_np((rank, float(loss.detach()), xl.grad.clone(), yl.grad.clone())))
    finally:
        dist.destroy_process_group()


def test_sharded_cloob_world2_gloo():
    """ShardedCLOOBLoss on 2 ranks against the single-process oracle (cloobLoss.py:10-47
    restated in oracle/) on the concatenated batch: same loss on every rank, gradients of the
    global loss w.r.t. the local shards."""
    from oracle import losses_oracle as O
    world, n, d = 2, 28, 24
    res = _run_world(_worker_cloob, world, (n, d))
    x, y = O.synth_embeddings(n * world, d, 9, correlated=True)
    ref, (rx, ry) = O.loss_and_grads(lambda a, b: 3.0 * O.cloob_loss(a, b, 30.0, 15.0), [x, y])
    for rank, loss, gx, gy in res:
        assert abs(3.0 * loss - float(ref)) < 1e-5 * abs(float(ref))
        sl = slice(rank * n, (rank + 1) * n)
        assert (gx - rx[sl]).norm() < 1e-4 * rx[sl].norm()
        assert (gy - ry[sl]).norm() < 1e-4 * ry[sl].norm()


def _worker_sampled(rank, world, port, q, eps, S, n, d, ls):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cpu_backend
        from multimodal_2025_b200.dist import ShardedClipLoss
        g = torch.Generator().manual_seed(31)
        x = torch.randn(S, n * world, d, generator=g) * 1.3
        y = x + 2.5 * torch.randn(S, n * world, d, generator=g)
        sl = slice(rank * n, (rank + 1) * n)
        xl = x[:, sl].clone().requires_grad_(True)
        yl = y[:, sl].clone().requires_grad_(True)
        lsp = torch.tensor(ls, requires_grad=True)
        crit = ShardedClipLoss(eps, backend=cpu_backend)
        loss = crit.forward_sampled(xl, yl, lsp)
        w = torch.arange(1, S + 1, dtype=torch.float32)      # distinct upstream gradients
        (loss * w).sum().backward()
        q.put(_np((rank, loss.detach().clone(), xl.grad.clone(), yl.grad.clone(), float(lsp.grad))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("eps,ls", [(0.1, 2.6592), (0.0, 4.5)])
def test_sharded_clip_forward_sampled_world2_gloo(eps, ls):
    """ShardedClipLoss.forward_sampled on 2 ranks against the oracle's forward_sampled
    (clipLoss.py:24-79 restated) on the concatenated [S, N, E] samples."""
    from oracle import losses_oracle as O
    world, S, n, d = 2, 3, 20, 16
    res = _run_world(_worker_sampled, world, (eps, S, n, d, ls))
    g = torch.Generator().manual_seed(31)
    x = torch.randn(S, n * world, d, generator=g) * 1.3
    y = x + 2.5 * torch.randn(S, n * world, d, generator=g)
    w = torch.arange(1, S + 1, dtype=torch.float32)
    ref = O.clip_loss_sampled(x, y, torch.tensor(ls), eps)
    _, (rx, ry, rl) = O.loss_and_grads(
        lambda a, b, c: (O.clip_loss_sampled(a, b, c, eps) * w).sum(), [x, y, torch.tensor(ls)])
    for rank, loss, gx, gy, gl in res:
        assert (loss - ref).abs().max() < 1e-5 * ref.abs().max()
        sl = slice(rank * n, (rank + 1) * n)
        assert (gx - rx[:, sl]).norm() < 1e-4 * rx[:, sl].norm()
        assert (gy - ry[:, sl]).norm() < 1e-4 * ry[:, sl].norm()
        assert abs(gl - float(rl)) <= 1e-4 * max(abs(float(rl)), 1e-6)


def _vclip_inputs(N, d):
    g = torch.Generator().manual_seed(41)
    loc_i = torch.nn.functional.normalize(torch.randn(N, d, generator=g), dim=-1)
    loc_t = torch.nn.functional.normalize(loc_i + 0.8 * torch.randn(N, d, generator=g), dim=-1)
    k_i = torch.rand(N, generator=g) * 90 + 10
    k_t = torch.rand(N, generator=g) * 90 + 10
    return loc_i, loc_t, k_i, k_t


def _worker_vclip(rank, world, port, q, n, d):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cpu_backend
        from multimodal_2025_b200.dist import make_sharded_vclip_loss
        from multimodal_2025_b200.losses.distributions.power_spherical import PowerSpherical
        loc_i, loc_t, k_i, k_t = _vclip_inputs(n * world, d)
        sl = slice(rank * n, (rank + 1) * n)
        li = loc_i[sl].clone().requires_grad_(True)
        lt = loc_t[sl].clone().requires_grad_(True)
        ki = k_i[sl].clone().requires_grad_(True)
        kt = k_t[sl].clone().requires_grad_(True)
        crit = make_sharded_vclip_loss(backend=cpu_backend, kl_weight=0.7, var_reg_weight=0.2)
        out = crit(PowerSpherical(li, ki), PowerSpherical(lt, kt), 1.0 / ki, 1.0 / kt,
                   torch.tensor(2.6592), is_train=False)
        out["total_loss"].backward()
        q.put(_np((rank, {k: float(v.detach()) for k, v in out.items()}, li.grad.clone(),
               ki.grad.clone(), kt.grad.clone())))
    finally:
        dist.destroy_process_group()


def test_sharded_vclip_eval_path_world2_gloo():
    """make_sharded_vclip_loss on the no-sampling branch (vclipLoss.py:104-108: distribution
    means -> ClipLoss.forward) + global KL / var-reg means, against the same formula evaluated
    on the concatenated batch with the oracle's CLIP loss (the sampling branch needs the CUDA
    sampler kernels; its sharded CLIP term is covered by the forward_sampled test above)."""
    from multimodal_2025_b200.losses.distributions.power_spherical import (HypersphericalUniform,
                                                                           PowerSpherical)
    from oracle import losses_oracle as O
    world, n, d = 2, 18, 16
    res = _run_world(_worker_vclip, world, (n, d))
    loc_i, loc_t, k_i, k_t = [t.clone().requires_grad_(True) for t in _vclip_inputs(n * world, d)]
    clip = O.clip_loss(loc_i, loc_t, torch.tensor(2.6592), 0.1, True)
    kl = lambda loc, k: torch.distributions.kl_divergence(  # noqa: E731
        PowerSpherical(loc, k), HypersphericalUniform(dim=d)).mean()
    kli, klt = kl(loc_i, k_i), kl(loc_t, k_t)
    var = torch.mean((torch.log(1.0 / k_i + 1e-8) - torch.log(1.0 / k_t + 1e-8)) ** 2)
    total = 1.0 * clip + 0.5 * 0.7 * (kli + klt) + 0.2 * var
    total.backward()
    for rank, out, gli, gki, gkt in res:
        sl = slice(rank * n, (rank + 1) * n)
        for key, want in (("total_loss", total), ("clip_loss", clip), ("image_kl_loss", kli),
                          ("text_kl_loss", klt), ("var_reg", var)):
            assert abs(out[key] - float(want.detach())) < 2e-5 * abs(float(want.detach())), key
        assert (gli - loc_i.grad[sl]).norm() < 1e-4 * loc_i.grad[sl].norm()
        assert (gki - k_i.grad[sl]).norm() < 1e-4 * k_i.grad[sl].norm()
        assert (gkt - k_t.grad[sl]).norm() < 1e-4 * k_t.grad[sl].norm()


def _worker_flags(rank, world, port, q, n, d):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import torch.nn.functional as F
        from torch.distributed.nn import all_gather as ag_with_grad
        import cpu_backend
        from multimodal_2025_b200.losses.align_clip_loss import ClipLoss
        g = torch.Generator().manual_seed(11)
        X = F.normalize(torch.randn(n * world, d, generator=g), dim=-1)
        Y = F.normalize(X + 0.5 * torch.randn(n * world, d, generator=g), dim=-1)
        sl = slice(rank * n, (rank + 1) * n)
        out = {}
        for gwg in (False, True):
            # the reference's formula (loss.py:20-64, 103-130) written with the same collectives
            x = X[sl].clone().requires_grad_(True)
            y = Y[sl].clone().requires_grad_(True)
            if gwg:
                ax = torch.cat(ag_with_grad(x), dim=0)
                ay = torch.cat(ag_with_grad(y), dim=0)
            else:
                gx = [torch.zeros_like(x) for _ in range(world)]
                gy = [torch.zeros_like(y) for _ in range(world)]
                dist.all_gather(gx, x.detach())
                dist.all_gather(gy, y.detach())
                gx[rank], gy[rank] = x, y
                ax, ay = torch.cat(gx), torch.cat(gy)
            z = 14.285 * ax @ ay.t()
            lab = torch.arange(n * world)
            ref = (F.cross_entropy(z, lab) + F.cross_entropy(z.t(), lab)) / 2
            ref.backward()
            crit = ClipLoss(gather_with_grad=gwg, rank=rank, world_size=world)
            crit._backend = cpu_backend
            xo = X[sl].clone().requires_grad_(True)
            yo = Y[sl].clone().requires_grad_(True)
            loss = crit(xo, yo, 14.285)
            loss.backward()
            out[gwg] = (ref.detach(), x.grad, y.grad, loss.detach(), xo.grad, yo.grad)
        # local_loss=True: each rank's own loss (loss.py:108-110, labels + n * rank :94-95)
        for gwg in (False, True):
            x = X[sl].clone().requires_grad_(True)
            y = Y[sl].clone().requires_grad_(True)
            if gwg:
                ax = torch.cat(ag_with_grad(x), dim=0)
                ay = torch.cat(ag_with_grad(y), dim=0)
            else:
                gx = [torch.zeros_like(x) for _ in range(world)]
                gy = [torch.zeros_like(y) for _ in range(world)]
                dist.all_gather(gx, x.detach())
                dist.all_gather(gy, y.detach())
                ax, ay = torch.cat(gx), torch.cat(gy)
            lab = torch.arange(n) + n * rank
            ref = (F.cross_entropy(14.285 * x @ ay.t(), lab) +
                   F.cross_entropy(14.285 * y @ ax.t(), lab)) / 2
            (ref * (1.0 + rank)).backward()     # a different upstream gradient on every rank
            crit = ClipLoss(local_loss=True, gather_with_grad=gwg, rank=rank, world_size=world)
            crit._backend = cpu_backend
            xo = X[sl].clone().requires_grad_(True)
            yo = Y[sl].clone().requires_grad_(True)
            loss = crit(xo, yo, 14.285)
            (loss * (1.0 + rank)).backward()
            out[("local", gwg)] = (ref.detach(), x.grad, y.grad, loss.detach(), xo.grad, yo.grad)
        errs = []
        for kw in (dict(use_horovod=True),):
            try:
                c = ClipLoss(rank=rank, world_size=world, **kw)
                c._backend = cpu_backend
                c(X[sl], Y[sl], 14.285)
                errs.append(None)
            except NotImplementedError as e:
                errs.append(str(e))
        try:
            ClipLoss(rank=rank + 1, world_size=world)(X[sl], Y[sl], 14.285)
            errs.append(None)
        except ValueError as e:
            errs.append(str(e))
        try:
            from multimodal_2025_b200.losses.align_clip_loss import ClipInModalityLoss
            c = ClipInModalityLoss(local_loss=True, rank=rank, world_size=world)
            c(X[sl], Y[sl], 14.285)
            errs.append(None)
        except NotImplementedError as e:
            errs.append(str(e))
        q.put(_np((rank, out, errs)))
    finally:
        dist.destroy_process_group()


def test_dropin_cliploss_distributed_flags_world2_gloo():
    """the drop-in open_clip ClipLoss maps gather_with_grad onto the gradient
    convention of the reference (own-slot gradient vs sum over ranks), computes the per-rank
    loss for local_loss=True (both gradient conventions, distinct upstream gradients), and
    refuses horovod / a rank that does not match the group instead of silently ignoring them."""
    res = _run_world(_worker_flags, 2, (12, 16))
    for rank, out, errs in res:
        for gwg, (ref, rx, ry, loss, gx, gy) in out.items():
            assert abs(float(loss) - float(ref)) < 1e-5 * abs(float(ref)), gwg
            assert (gx - rx).norm() < 1e-4 * rx.norm(), gwg
            assert (gy - ry).norm() < 1e-4 * ry.norm(), gwg
        assert all(e is not None for e in errs), errs


@pytest.mark.parametrize("need_scale", [True, False])
def test_cached_negatives_host_logic_with_standin(need_scale):
    """SURVEY §8f rank 4 (main/train.py:139-157): micro-step j of the accumulation loop — loss of
    the spliced batch, gradients of the LIVE rows only — against autograd through the
    reference formula on the spliced tensors; dense stand-in ops (the GPU test runs the kernels)."""
    import torch.nn.functional as F
    import cpu_backend
    from multimodal_2025_b200.losses.negatives_cache import cached_negatives_clip_loss
    g = torch.Generator().manual_seed(2)
    k, b, d, j = 4, 6, 16, 2
    ci = [F.normalize(torch.randn(b, d, generator=g), dim=-1) for _ in range(k)]
    ct = [F.normalize(torch.randn(b, d, generator=g), dim=-1) for _ in range(k)]
    li = F.normalize(ci[j] + 0.05 * torch.randn(b, d, generator=g), dim=-1)
    lt = F.normalize(ct[j] + 0.05 * torch.randn(b, d, generator=g), dim=-1)
    # reference: loss(**inputs) with inputs = cat(accumulated[:j] + [live] + accumulated[j+1:])
    a = li.clone().requires_grad_(True)
    t = lt.clone().requires_grad_(True)
    s = torch.tensor(14.285, requires_grad=need_scale)
    xi = torch.cat(ci[:j] + [a] + ci[j + 1:])
    xt = torch.cat(ct[:j] + [t] + ct[j + 1:])
    z = s * xi @ xt.T
    lab = torch.arange(k * b)
    ref = (F.cross_entropy(z, lab) + F.cross_entropy(z.T, lab)) / 2
    ref.backward()
    a2 = li.clone().requires_grad_(True)
    t2 = lt.clone().requires_grad_(True)
    s2 = torch.tensor(14.285, requires_grad=need_scale)
    loss = cached_negatives_clip_loss(a2, t2, s2, ci, ct, j, backend=cpu_backend)
    loss.backward()
    assert abs(float(loss) - float(ref)) < 1e-5 * abs(float(ref))
    assert (a2.grad - a.grad).norm() < 1e-4 * a.grad.norm()
    assert (t2.grad - t.grad).norm() < 1e-4 * t.grad.norm()
    if need_scale:
        assert abs(float(s2.grad) - float(s.grad)) < 1e-4 * abs(float(s.grad))


def _worker_seed(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from multimodal_2025_b200.losses.distributions._fn import fresh_seed
        torch.manual_seed(1234)          # every rank seeds identically (main/train.py does)
        a = fresh_seed()
        torch.manual_seed(1234)
        b = fresh_seed()
        q.put((rank, a, b))
    finally:
        dist.destroy_process_group()


def test_philox_key_differs_across_ranks_and_is_reproducible():
    """identical torch.manual_seed on every rank must not give the ranks identical
    sampler noise; the key stays reproducible per rank."""
    res = _run_world(_worker_seed, 2, ())
    (r0, a0, b0), (r1, a1, b1) = res
    assert a0 == b0 and a1 == b1 and a0 != a1
    assert 0 <= a1 < 2 ** 63
