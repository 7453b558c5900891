"""numpy restatement of handyrl/losses.py (TEST INFRASTRUCTURE, see oracle/__init__.py).

fp32 throughout, vectorised over the batch, sequential over time exactly like
the reference's Python loop.  Scalars follow torch's rule for `python_float *
fp32_tensor`: the double is rounded to fp32 once; `(1 - lmb)` and `gamma * lmb`
are formed in double first (losses.py:24,36,51).
"""
import numpy as np

f32 = np.float32


def monte_carlo(values, returns):
    # losses.py:16-17
    return returns, (returns - values).astype(f32)


def temporal_difference(values, returns, rewards, lmb, gamma):
    # losses.py:20-28
    T = values.shape[1]
    tv = [None] * T
    tv[T - 1] = np.broadcast_to(returns[:, -1], np.broadcast_shapes(returns[:, -1].shape, values[:, 0].shape)).astype(f32)
    oml, l, g = f32(1 - lmb), f32(lmb), f32(gamma)
    for i in range(T - 2, -1, -1):
        reward = rewards[:, i] if rewards is not None else f32(0)
        tv[i] = (reward + g * (oml * values[:, i + 1] + l * tv[i + 1])).astype(f32)
    tv = np.stack(np.broadcast_arrays(*tv), axis=1)
    return tv, (tv - values).astype(f32)


def upgo(values, returns, rewards, lmb, gamma):
    # losses.py:31-40
    T = values.shape[1]
    tv = [None] * T
    tv[T - 1] = np.broadcast_to(returns[:, -1], np.broadcast_shapes(returns[:, -1].shape, values[:, 0].shape)).astype(f32)
    oml, l, g = f32(1 - lmb), f32(lmb), f32(gamma)
    for i in range(T - 2, -1, -1):
        value = values[:, i + 1]
        reward = rewards[:, i] if rewards is not None else f32(0)
        tv[i] = (reward + g * np.maximum(value, oml * value + l * tv[i + 1])).astype(f32)
    tv = np.stack(np.broadcast_arrays(*tv), axis=1)
    return tv, (tv - values).astype(f32)


def vtrace(values, returns, rewards, lmb, gamma, rhos, cs):
    # losses.py:43-58
    T = values.shape[1]
    rewards = rewards if rewards is not None else f32(0)
    g, gl = f32(gamma), f32(gamma * lmb)
    last = np.broadcast_to(returns[:, -1:], values[:, -1:].shape)
    values_t_plus_1 = np.concatenate([values[:, 1:], last], axis=1)
    deltas = (rhos * ((rewards + g * values_t_plus_1).astype(f32) - values)).astype(f32)
    xs = [None] * T
    xs[T - 1] = deltas[:, -1]
    for i in range(T - 2, -1, -1):
        xs[i] = (deltas[:, i] + (gl * cs[:, i]).astype(f32) * xs[i + 1]).astype(f32)
    xs = np.stack(np.broadcast_arrays(*xs), axis=1)
    vs = (xs + values).astype(f32)
    vs_t_plus_1 = np.concatenate([vs[:, 1:], np.broadcast_to(returns[:, -1:], vs[:, -1:].shape)], axis=1)
    advantages = ((rewards + g * vs_t_plus_1).astype(f32) - values).astype(f32)
    return vs, advantages


def compute_target(algorithm, values, returns, rewards, lmb, gamma, rhos, cs):
    # losses.py:61-74
    if values is None:
        return None, 0
    if algorithm == 'MC':
        return monte_carlo(values, returns)
    elif algorithm == 'TD':
        return temporal_difference(values, returns, rewards, lmb, gamma)
    elif algorithm == 'UPGO':
        return upgo(values, returns, rewards, lmb, gamma)
    elif algorithm == 'VTRACE':
        return vtrace(values, returns, rewards, lmb, gamma, rhos, cs)
    else:
        print('No algorithm named %s' % algorithm)
