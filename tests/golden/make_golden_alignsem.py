"""Golden vectors for the AlignCLIP SEMANTIC branch (AlignCLIP/align_clip/loss.py:226-255), made by
RUNNING THE UNMODIFIED REFERENCE CLASS on the CPU.

The branch calls `logits.get_device()` (-1 on CPU) and then `torch.eye(n).to(-1)`, which raises
(loss.py:233,237).  For the duration of the reference call `torch.Tensor.get_device` is patched
to return the string 'cpu'; nothing in the reference module is edited, and every arithmetic
line executes as written.  Also runs `losses/alignCLIPLoss.py::AlignCLIPLoss` (the wrapper the
north star names) with the AlignCLIP package stubbed to the loss module loaded by path.

Run in the build container only:   python tests/golden/make_golden_alignsem.py
Writes tests/golden/alignclip_sem.npz.
"""
import contextlib
import importlib.util
import os
import sys
import types

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, REF, load_by_path, run, synth  # noqa: E402


@contextlib.contextmanager
def cpu_get_device():
    orig = torch.Tensor.get_device
    torch.Tensor.get_device = lambda self: "cpu"
    try:
        yield
    finally:
        torch.Tensor.get_device = orig


def main():
    torch.set_num_threads(4)
    align_mod = load_by_path("ref_alignloss", "AlignCLIP/align_clip/loss.py")
    out = {}
    cases = {
        # tag: (n, d, d_sem, seed, scale, separate_text, separate_image, alpha, beta)
        "sem_text_n64": (64, 32, 16, 31, 14.285, True, False, 1.0, 0.5),      # AlignCLIPLoss defaults
        "sem_both_n64": (64, 32, 16, 32, 14.285, True, True, 1.0, 0.5),       # train_alignCLIP.sh:18
        "sem_image_n96": (96, 48, 24, 33, 20.0, False, True, 0.7, 0.25),
        "sem_text_n160": (160, 64, 40, 34, 100.0, True, False, 1.0, 0.5),     # scale at the clamp
        "sem_both_n200_ragged": (200, 72, 56, 35, 8.0, True, True, 1.0, 0.5),
    }
    for tag, (n, d, ds, seed, sc, st, si, alpha, beta) in cases.items():
        x, y = synth(n, d, seed, True)
        g = torch.Generator().manual_seed(seed + 1000)
        # SBERT-like: un-normalised, correlated in groups so that 1 - S S^T is far from 1
        base = torch.randn(max(n // 8, 1), ds, generator=g)
        sem = base[torch.arange(n) % base.shape[0]] * 2.0 + torch.randn(n, ds, generator=g)
        s = torch.tensor(sc)
        crit = align_mod.ClipInModalityLoss(alpha=alpha, beta=beta, nl_semantic_supervision=True,
                                            separate_text=st, separate_image=si)
        with cpu_get_device():
            loss, (gx, gy, gs) = run(lambda a, b, c: crit(a, b, c, semantic_features=sem), x, y, s)
            d_out = crit(x, y, s, output_dict=True, semantic_features=sem)
        out[tag] = dict(x=x.numpy(), y=y.numpy(), sem=sem.numpy(), s=np.float32(sc),
                        flags=np.array([st, si], dtype=np.int64),
                        ab=np.array([alpha, beta], dtype=np.float32),
                        loss=loss, gx=gx, gy=gy, gs=gs,
                        clip_loss=d_out["clip_loss"].numpy(),
                        inmod_loss=d_out["inModality_loss"].numpy())

    # the wrapper itself: losses/alignCLIPLoss.py imports `AlignCLIP.align_clip.loss`, whose package
    # __init__ needs ftfy; expose the already-loaded loss module under that dotted name
    for name in ("AlignCLIP", "AlignCLIP.align_clip"):
        pkg = types.ModuleType(name)
        pkg.__path__ = []
        sys.modules[name] = pkg
    sys.modules["AlignCLIP.align_clip.loss"] = align_mod
    wrap_mod = load_by_path("ref_alignCLIPLoss", "losses/alignCLIPLoss.py")
    c = out["sem_text_n64"]
    w = wrap_mod.AlignCLIPLoss()
    sem = torch.from_numpy(c["sem"])
    with cpu_get_device():
        loss, (gx, gy, gs) = run(lambda a, b, s_: w.forward(a, b, s_, semantic_features=sem),
                                 torch.from_numpy(c["x"]), torch.from_numpy(c["y"]),
                                 torch.tensor(float(c["s"])))
    assert np.array_equal(loss, c["loss"]) and np.array_equal(gx, c["gx"])
    out["wrapper_n64"] = dict(loss=loss, gx=gx, gy=gy, gs=gs)
    np.savez_compressed(os.path.join(OUT, "alignclip_sem.npz"),
                        **{f"{t}/{k}": v for t, c in out.items() for k, v in c.items()})
    print("written", os.path.join(OUT, "alignclip_sem.npz"),
          {t: float(c["loss"]) for t, c in out.items()})


if __name__ == "__main__":
    main()
