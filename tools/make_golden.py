"""Generate tests/golden/*.npz by running the REFERENCE itself (imported read-only
from /root/reference).  Runs only in the build container; the GPU box and the
tests read just the committed fixtures.

    PYTHONDONTWRITEBYTECODE=1 python tools/make_golden.py [targets|ttt|geister|batch|all]
"""
import os
import sys

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("HANDYRL_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

import numpy as np  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def gen_targets():
    import torch
    from handyrl.losses import compute_target  # the reference
    from tests.synth import target_inputs, TARGET_GOLDEN_CASES

    out = {}
    for ci, (B, T, P, head, rpp) in enumerate(TARGET_GOLDEN_CASES):
        d = target_inputs(B, T, P, head, rpp, seed=100 + ci)
        tt = {k: (torch.from_numpy(v) if isinstance(v, np.ndarray) else v) for k, v in d.items()}
        for k in ('values', 'returns', 'rewards', 'rhos', 'cs'):
            if d[k] is not None:
                out['c%d_%s' % (ci, k)] = d[k]
        for mode in ('MC', 'TD', 'UPGO', 'VTRACE'):
            tg, adv = compute_target(mode, tt['values'], tt['returns'], tt['rewards'], d['lmb'], d['gamma'],
                                     tt['rhos'], tt['cs'])
            out['c%d_%s_targets' % (ci, mode)] = tg.numpy()
            out['c%d_%s_advantages' % (ci, mode)] = adv.numpy()
    np.savez_compressed(os.path.join(GOLD, "targets.npz"), **out)
    print("targets.npz:", len(out), "arrays")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    os.makedirs(GOLD, exist_ok=True)
    table = {k[4:]: v for k, v in list(globals().items()) if k.startswith("gen_")}
    for name, fn in table.items():
        if what in (name, "all"):
            fn()
