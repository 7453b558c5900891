"""numpy restatement of `make_batch` (handyrl/train.py:33-133) — TEST INFRASTRUCTURE
(see oracle/__init__.py).  Works on reference-format windows (`ep_minimum` dicts of
train.py:296-301 whose `moment` blocks are bz2-pickled lists of moments) and returns numpy
arrays keyed like the reference's dict (dict observations are flattened to
'observation.<key>').  Pinned against outputs of the reference itself
(tests/golden/batch_cases.pkl, tools/make_golden.py: gen_batch).
"""
import bz2
import pickle
import random

import numpy as np


def _leaves(obs):
    """observation -> ordered {name: array}; plain arrays get the empty name"""
    if isinstance(obs, dict):
        return {k: np.asarray(v) for k, v in obs.items()}
    return {'': np.asarray(obs)}


def _window_moments(ep):
    moments = []
    for blob in ep['moment']:                                   # train.py:54
        moments.extend(pickle.loads(bz2.decompress(blob)))
    return moments[ep['start'] - ep['base']:ep['end'] - ep['base']]  # train.py:55


def episode_arrays(ep, args, rng=random):
    """Per-episode (T', P, ...) arrays before padding (train.py:56-89)."""
    ms = _window_moments(ep)
    players = list(ms[0]['observation'].keys())                 # train.py:56
    if not args['turn_based_training']:
        players = [rng.choice(players)]                         # train.py:57-58
    first_turn = ms[0]['turn'][0]
    obs_tmpl = {k: np.zeros_like(v) for k, v in _leaves(ms[0]['observation'][first_turn]).items()}
    pol_tmpl = np.zeros_like(ms[0]['policy'][first_turn])
    turn_only = args['turn_based_training'] and not args['observation']   # train.py:64
    rows = []
    for m in ms:
        cols = [m['turn'][0]] if turn_only else players
        r = {'obs': [], 'policy': [], 'action': [], 'amask': []}
        for p in cols:
            o = m['observation'][p]
            r['obs'].append(obs_tmpl if o is None else _leaves(o))
            r['policy'].append(pol_tmpl if m['policy'][p] is None else m['policy'][p])
            r['action'].append(0 if m['action'][p] is None else m['action'][p])
            r['amask'].append(pol_tmpl + 1e32 if m['action_mask'][p] is None else m['action_mask'][p])
        rows.append(r)
    n = len(ms)
    out = {}
    for k in obs_tmpl:
        out['observation' + ('.' + k if k else '')] = np.array([[c[k] for c in r['obs']] for r in rows])
    out['policy'] = np.array([r['policy'] for r in rows])
    out['action'] = np.array([r['action'] for r in rows], dtype=np.int64)[..., np.newaxis]
    out['action_mask'] = np.array([r['amask'] for r in rows])

    def per_player(key, dtype=np.float32):                      # train.py:80-82
        return np.array([[[0] if m[key][p] is None else m[key][p] for p in players] for m in ms],
                        dtype=dtype).reshape(n, len(players), -1)

    out['value'] = per_player('value')
    out['reward'] = per_player('reward')
    out['return'] = per_player('return')
    out['outcome'] = np.array([ep['outcome'][p] for p in players], dtype=np.float32).reshape(1, len(players), -1)
    out['episode_mask'] = np.ones((n, 1, 1), dtype=np.float32)
    out['turn_mask'] = np.array([[[m['policy'][p] is not None] for p in players] for m in ms], dtype=np.float32)
    out['observation_mask'] = np.array([[[m['value'][p] is not None] for p in players] for m in ms],
                                       dtype=np.float32)
    out['progress'] = np.arange(ep['start'], ep['end'], dtype=np.float32)[..., np.newaxis] / ep['total']
    return out


# constant used to fill rows t >= window length (train.py:92-104); 'value' is filled with the outcome
PAD = {'policy': 0, 'action': 0, 'reward': 0, 'return': 0, 'episode_mask': 0, 'turn_mask': 0,
       'observation_mask': 0, 'action_mask': 1e32, 'progress': 1}


def make_batch(episodes, args, rng=random):
    T = args['forward_steps']
    per_ep = []
    for ep in episodes:
        a = episode_arrays(ep, args, rng)
        n = a['turn_mask'].shape[0]
        if n < T:
            pad = T - n
            for k in list(a):
                if k == 'outcome':
                    continue
                if k == 'value':
                    a[k] = np.concatenate([a[k], np.tile(a['outcome'], [pad, 1, 1])])       # train.py:96
                    continue
                fill = 0 if k.startswith('observation') and k != 'observation_mask' else PAD[k]
                width = [(0, pad)] + [(0, 0)] * (a[k].ndim - 1)
                a[k] = np.pad(a[k], width, 'constant', constant_values=fill)
        per_ep.append(a)
    return {k: np.array([a[k] for a in per_ep]) for k in per_ep[0]}
