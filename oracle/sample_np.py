"""Restatement of the action draw of handyrl/generation.py:48-53 for a given uniform number —
TEST INFRASTRUCTURE (see oracle/__init__.py).

    p = logits - action_mask;  action = random.choices(legal_actions, weights=softmax(p[legal_actions]))[0]

`softmax` is util.py:61-63 in fp32.  `random.choices` (CPython Lib/random.py) does
`cum_weights = list(accumulate(weights))`, `total = cum_weights[-1] + 0.0` and returns
`population[bisect(cum_weights, random() * total, 0, n - 1)]`.  The weights are numpy float32 scalars, so the
running sum is rounded to fp32 at every step; `random() * total` is a double under numpy 1.x scalar rules and an
fp32 under numpy >= 2 (NEP 50) - the two differ by < 1e-7 relative, far below what separates two actions except
within rounding of a boundary.  `u` stands for that `random()`.

Pinned by tests/golden/sampler_cases.npz: 2,200 states of reference games with the `random()` value that
`random.choices` consumed and the action it returned (tools/make_golden.py gen_sampler)."""
from bisect import bisect
from itertools import accumulate

import numpy as np


def softmax(x):
    x = np.exp(x - np.max(x, axis=-1))
    return x / x.sum(axis=-1)


def choose(legal_actions, logits, u):
    p = np.asarray(logits, dtype=np.float32)[list(legal_actions)]
    w = softmax(p)
    cum = list(accumulate(w))                    # np.float32 scalars: fp32 running sum, like the reference
    total = cum[-1] + 0.0
    k = bisect(cum, u * total, 0, len(cum) - 1)
    return int(legal_actions[k]), [float(c) for c in cum]
