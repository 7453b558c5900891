"""Restatement of the action draw of handyrl/generation.py:48-53 for a given uniform number —
TEST INFRASTRUCTURE (see oracle/__init__.py).

    p = logits - action_mask;  action = random.choices(legal_actions, weights=softmax(p[legal_actions]))[0]

`softmax` is util.py:61-63 in fp32; `random.choices` accumulates the weights as Python floats and returns
`population[bisect(cum_weights, random() * total, 0, n - 1)]`.  `u` stands for that `random()`."""
from bisect import bisect
from itertools import accumulate

import numpy as np


def softmax(x):
    x = np.exp(x - np.max(x, axis=-1))
    return x / x.sum(axis=-1)


def choose(legal_actions, logits, u):
    p = np.asarray(logits, dtype=np.float32)[list(legal_actions)]
    w = softmax(p)
    cum = list(accumulate(float(x) for x in w))
    k = bisect(cum, u * cum[-1], 0, len(cum) - 1)
    return int(legal_actions[k]), cum
