"""Seeded synthetic inputs shared by the parity tests, bench.py and tools/make_golden.py.

compute_target inputs follow SURVEY.md §8 D4.
"""
import numpy as np


def target_inputs(B, T, P, head, rho_per_player=False, seed=0, gamma=None):
    """head = 'value'  -> returns (B,1,P,1) in {-1,0,1}, rewards None, gamma 1.0
       head = 'return' -> rewards -0.01, returns (B,T,P,1) discounted sums, gamma 0.8"""
    rng = np.random.RandomState(seed)
    v = np.tanh(rng.standard_normal((B, T, P, 1))).astype(np.float32)
    rho_raw = np.exp(rng.standard_normal((B, T, P if rho_per_player else 1, 1)) * 0.5)
    rho = np.clip(rho_raw, 0, 1).astype(np.float32)
    c = rho.copy()
    if head == 'value':
        returns = rng.randint(-1, 2, size=(B, 1, P, 1)).astype(np.float32)
        rewards = None
        g = 1.0 if gamma is None else gamma
    else:
        g = 0.8 if gamma is None else gamma
        rewards = np.full((B, T, P, 1), -0.01, dtype=np.float32)
        rewards = (rewards * (1 + 0.1 * rng.standard_normal((B, T, P, 1)))).astype(np.float32)
        returns = np.zeros((B, T, P, 1), dtype=np.float32)
        acc = np.zeros((B, P, 1), dtype=np.float32)
        for t in range(T - 1, -1, -1):
            acc = (rewards[:, t] + np.float32(g) * acc).astype(np.float32)
            returns[:, t] = acc
    return dict(values=v, returns=returns, rewards=rewards, rhos=rho, cs=c, lmb=0.7, gamma=g)


TARGET_GOLDEN_CASES = [
    # (B, T, P, head, rho_per_player)
    (5, 1, 2, 'value', False),
    (7, 2, 2, 'return', False),
    (64, 16, 2, 'value', False),
    (64, 16, 2, 'return', False),
    (33, 32, 2, 'return', True),
    (8, 200, 2, 'value', False),
    (3, 1024, 2, 'return', False),
    (16, 16, 1, 'value', False),
    (4, 37, 3, 'return', True),
    (9, 35, 2, 'value', False),
]
