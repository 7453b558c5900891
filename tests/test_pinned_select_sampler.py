"""A11 / F1 pinned to reference output (tests/golden/select_cases.pkl, sampler_cases.npz, written by
tools/make_golden.py from the imported reference):

* `handyrl_b200.train.select_episode` reproduces `Batcher.select_episode` (handyrl/train.py:284-302) draw for
  draw under the same Mersenne-Twister seed (episode, start, end, base, total, shipped blocks);
* the window law of the host-C / device selection routine (counter RNG, so not the same stream) agrees with the
  reference's draws in distribution and with the rule's exact law;
* the action draw (handyrl/generation.py:48-53 = softmax over the legal logits + random.choices) of the numpy
  restatement and of `b200rl_env_sample_record`, fed the `random()` number the reference consumed, returns the
  reference's action on every recorded state.
"""
import ctypes
import os
import pickle
import random

import numpy as np
import pytest

from oracle import sample_np


@pytest.fixture(scope="module")
def select_cases(golden_dir):
    with open(os.path.join(golden_dir, "select_cases.pkl"), "rb") as f:
        return pickle.load(f)


@pytest.fixture(scope="module")
def sampler_cases(golden_dir):
    z = np.load(os.path.join(golden_dir, "sampler_cases.npz"))
    out = {}
    for k in z.files:
        e, f = k.split('.', 1)
        out.setdefault(e, {})[f] = z[k]
    return out


def _episodes(c):
    cs = c['args']['compress_steps']
    return [{'args': {'id': i}, 'outcome': {0: 1, 1: -1}, 'steps': st,
             'moment': [(i, b) for b in range((st + cs - 1) // cs)]} for i, st in enumerate(c['steps'])]


@pytest.mark.parametrize("name", ['short_buffer', 'steep', 'overfull', 'one_step_windows', 'long_window'])
def test_select_episode_reproduces_reference_draws(select_cases, name):
    from handyrl_b200.train import select_episode
    c = select_cases[name]
    eps = _episodes(c)
    rng = random.Random(c['seed'])
    for row in c['rows']:
        w = select_episode(eps, c['args'], rng)
        got = (w['args']['id'], w['start'], w['end'], w['base'], w['total'], w['moment'][0][1], len(w['moment']))
        assert got == tuple(int(x) for x in row)


def _accept_law(L, M):
    """P(episode index) of the rejection loop of train.py:285-289"""
    n = min(L, M)
    acc = np.clip(1.0 - (L - 1 - np.arange(n)) / M, 0.0, 1.0)
    return acc / acc.sum()


@pytest.mark.parametrize("name", ['short_buffer', 'steep', 'overfull'])
def test_host_c_selection_law_matches_reference(select_cases, name):
    """b200rl_select_windows_host shares one __host__ __device__ routine with the device kernel
    (tests/test_select_windows.py checks them bit for bit).  Here its law is pinned to the reference."""
    from handyrl_b200 import _lib
    c = select_cases[name]
    steps = np.asarray(c['steps'])
    L, M, T = len(steps), c['args']['maximum_episodes'], c['args']['forward_steps']
    E = 512
    tab = np.zeros((E, 4), np.int32)
    end_it = np.cumsum(steps) + 1000                    # increasing, every episode alive (D huge)
    for i in range(L):
        tab[i] = (i, int(steps[i]) | (1 << 16), int(end_it[i]), 0)
    Bw = 200000
    out = np.zeros((Bw, 8), np.int32)
    _lib.call("b200rl_select_windows_host", ctypes.c_void_p(tab.ctypes.data), E, L, M, T, 1234, 0,
              int(end_it[-1]) + 1, 1 << 30, ctypes.c_void_p(out.ctypes.data), Bw, 2)
    idx, ln, st, tot = out[:, 0], out[:, 3], out[:, 4], out[:, 5]
    Lk = min(L, M)                                      # the trainer keeps the newest maximum_episodes (train.py:502)
    base = L - Lk
    assert idx.min() >= base and idx.max() < L
    assert np.array_equal(tot, steps[idx]) and np.array_equal(ln, np.minimum(st + T, tot) - st)
    assert np.all(st >= 0) and np.all(st <= np.maximum(0, tot - T))
    # (1) exact law of the rule
    law = _accept_law(Lk, M)
    cnt = np.bincount(idx - base, minlength=Lk).astype(np.float64)
    exp = law * Bw
    keep = exp >= 5
    chi2 = float((((cnt - exp) ** 2)[keep] / exp[keep]).sum())
    dof = int(keep.sum()) - 1
    assert chi2 < dof + 6 * np.sqrt(2 * dof), (chi2, dof)
    # (2) the reference's own draws (1,500 per case), coarse bins, two-sample chi-square
    if L <= M:      # with len(episodes) > maximum_episodes the reference only draws from the OLDEST M (train.py:286)
        ref_idx = c['rows'][:, 0]
        bins = np.linspace(0, L, 11).astype(int)
        a = np.histogram(idx, bins)[0].astype(np.float64)
        b = np.histogram(ref_idx, bins)[0].astype(np.float64)
        k1, k2 = np.sqrt(b.sum() / a.sum()), np.sqrt(a.sum() / b.sum())
        chi2 = float((((k1 * a - k2 * b) ** 2) / np.maximum(a + b, 1)).sum())
        assert chi2 < 9 + 6 * np.sqrt(18), chi2
    # (3) window start uniform over the candidates (train.py:291-292): mean start of the long episodes
    long_ep = tot > T + 20
    frac = st[long_ep] / (tot[long_ep] - T)
    assert abs(frac.mean() - 0.5) < 0.01


@pytest.mark.parametrize("env", ['tictactoe', 'geister'])
def test_sampler_oracle_reproduces_reference_actions(sampler_cases, env):
    c = sampler_cases[env]
    n = len(c['action'])
    assert n >= 700
    for i in range(n):
        legal = c['legal'][c['legal_off'][i]:c['legal_off'][i + 1]].tolist()
        a, _ = sample_np.choose(legal, c['logits'][i], float(c['u'][i]))
        assert a == int(c['action'][i]), i


@pytest.mark.gpu
@pytest.mark.parametrize("env,game", [('tictactoe', 'TicTacToe'), ('geister', 'Geister')])
def test_device_sampler_reproduces_reference_actions(sampler_cases, env, game):
    """Every game of the fixture in its own slot; each ply the device samples from the recorded logits with the
    recorded random() number, applies and records the action.  It must pick the reference's action, so the slot
    follows the reference game to its end without ever being told an action; the legal list it samples from is
    compared as well."""
    import torch
    from handyrl_b200.generation import BatchGenerator
    c = sampler_cases[env]
    off = c['tape_off']
    G = len(off) - 1
    L = int(np.diff(off).max())
    A = c['logits'].shape[1]
    row = {(int(g), int(p)): i for i, (g, p) in enumerate(zip(c['game'], c['ply']))}
    gen = BatchGenerator(game, G, seed=1, depth=L + 2)
    alive = np.ones(G, bool)
    checked = near = 0
    for k in range(L):
        logits = np.zeros((G, A), np.float32)
        u = np.full(G, 0.5)
        act_ref = -np.ones(G, np.int64)
        for g in range(G):
            i = row.get((g, k))
            if i is not None and alive[g]:
                logits[g], u[g], act_ref[g] = c['logits'][i], c['u'][i], c['action'][i]
        ids, cnt = gen.env.legal_list()
        ids, cnt = ids.cpu().numpy(), cnt.cpu().numpy()
        act, _ = gen.act(torch.from_numpy(logits).cuda(), u=torch.from_numpy(u).cuda())
        act = act.cpu().numpy()[:, 0]
        for g in np.nonzero(act_ref >= 0)[0]:
            i = row[(g, k)]
            legal = c['legal'][c['legal_off'][i]:c['legal_off'][i + 1]].tolist()
            assert ids[g, :cnt[g]].tolist() == legal
            checked += 1
            if act[g] != act_ref[g]:
                # CUDA expf and numpy's expf may differ in the last bit: only a draw within rounding of a
                # cumulative-weight boundary may flip; the slot then leaves the reference game
                _, cum = sample_np.choose(legal, c['logits'][i], float(c['u'][i]))
                x = float(c['u'][i]) * cum[-1]
                assert min(abs(x - cc) for cc in cum) < 1e-6 * cum[-1], (g, k)
                near += 1
                alive[g] = False
    assert checked >= 0.97 * len(c['action']) and near <= 2, (checked, near)
