"""make_batch for episodes in the reference wire format (bz2-pickled moment blocks,
handyrl/generation.py:79-86).

The blocks have to be decompressed and unpickled on the host (handyrl/train.py:54); what
moves to the GPU is everything after that: the per-episode windows of every key are
uploaded back to back (ragged) from pinned memory and padded + stacked into the
(B, T, P, ...) tensors by `b200rl_pad_stack` (train.py:92-122).  Works for any environment
whose observations are arrays or flat dicts of arrays, not only the built-in games.
"""
import bz2
import ctypes
import pickle
import random

import numpy as np
import torch

from . import _lib
from .engine import _ptr


def _obs_leaves(o):
    if isinstance(o, dict):
        return {k: np.asarray(v) for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return {i: np.asarray(v) for i, v in enumerate(o)}
    return {None: np.asarray(o)}


def _decode(ep):
    moments = []
    for blob in ep['moment']:
        moments += pickle.loads(bz2.decompress(blob))
    return moments[ep['start'] - ep['base']:ep['end'] - ep['base']]


def _episode_rows(ep, args):
    """ragged rows (T', ...) per key for one window, before padding (train.py:56-89)"""
    ms = _decode(ep)
    players = list(ms[0]['observation'].keys())
    if not args['turn_based_training']:
        players = [random.choice(players)]
    t0 = ms[0]['turn'][0]
    obs_zero = {k: np.zeros_like(v) for k, v in _obs_leaves(ms[0]['observation'][t0]).items()}
    p_zero = np.zeros_like(ms[0]['policy'][t0])
    turn_only = args['turn_based_training'] and not args['observation']
    n = len(ms)
    obs = {k: [] for k in obs_zero}
    pol, act, am = [], [], []
    for m in ms:
        cols = [m['turn'][0]] if turn_only else players
        for k in obs_zero:
            obs[k].append([obs_zero[k] if m['observation'][p] is None else _obs_leaves(m['observation'][p])[k]
                           for p in cols])
        pol.append([p_zero if m['policy'][p] is None else m['policy'][p] for p in cols])
        act.append([0 if m['action'][p] is None else m['action'][p] for p in cols])
        am.append([p_zero + 1e32 if m['action_mask'][p] is None else m['action_mask'][p] for p in cols])

    def pp(key):
        return np.array([[[0] if m[key][p] is None else m[key][p] for p in players] for m in ms],
                        dtype=np.float32).reshape(n, len(players), -1)

    rows = {('observation', k): np.array(v) for k, v in obs.items()}
    rows['policy'] = np.array(pol)
    rows['action'] = np.array(act, dtype=np.int64)[..., np.newaxis]
    rows['action_mask'] = np.array(am)
    rows['value'], rows['reward'], rows['return'] = pp('value'), pp('reward'), pp('return')
    rows['episode_mask'] = np.ones((n, 1, 1), dtype=np.float32)
    rows['turn_mask'] = np.array([[[m['policy'][p] is not None] for p in players] for m in ms], dtype=np.float32)
    rows['observation_mask'] = np.array([[[m['value'][p] is not None] for p in players] for m in ms],
                                        dtype=np.float32)
    rows['progress'] = np.arange(ep['start'], ep['end'], dtype=np.float32)[..., np.newaxis] / ep['total']
    outcome = np.array([ep['outcome'][p] for p in players], dtype=np.float32).reshape(1, len(players), -1)
    return rows, outcome


# fill of rows t >= window length (train.py:92-104)
_FILL = {'policy': 0.0, 'action': None, 'reward': 0.0, 'return': 0.0, 'episode_mask': 0.0, 'turn_mask': 0.0,
         'observation_mask': 0.0, 'action_mask': 1e32, 'progress': 1.0}


def make_batch_from_reference_episodes(episodes, args, device=None):
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    if dev.type != "cuda":
        raise _lib.B200RLError("make_batch runs on a CUDA device (no CPU fallback)")
    B, T = len(episodes), int(args['forward_steps'])
    per_ep, outcomes = zip(*[_episode_rows(ep, args) for ep in episodes])
    lens = [r['turn_mask'].shape[0] for r in per_ep]
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)).to(dev)
    outcome = np.stack(outcomes)                                     # (B, 1, P, 1)
    oc_dev = torch.from_numpy(outcome).to(dev)
    out = {}
    for key in per_ep[0]:
        rag = np.ascontiguousarray(np.concatenate([r[key] for r in per_ep]))   # (sum T', ...)
        tail, dt = rag.shape[1:], rag.dtype
        if key == 'value':
            fill_rows, fill = oc_dev, 0.0                            # value <- outcome (train.py:96)
        else:
            fill_rows = None
            fill = 0.0 if isinstance(key, tuple) else _FILL[key]
        if fill not in (None, 0.0) or fill_rows is not None:
            if dt != np.float32:                                     # the fill is a 32-bit float pattern
                raise _lib.B200RLError("make_batch: '%s' must be float32 (got %s)" % (key, dt))
        # The kernel moves rows as 4-byte words whatever the element type (uint8 images, bool planes, float64
        # vectors ... - the reference's to_torch keeps the numpy dtype, model.py:16-19): rows are handed over as raw
        # bytes, padded on the host to a multiple of 4 when needed.  Observation padding is all-zero bytes.
        row_bytes = int(np.prod(tail, dtype=np.int64)) * dt.itemsize
        raw = rag.view(np.uint8).reshape(rag.shape[0], row_bytes)
        pad = (-row_bytes) % 4
        if pad:
            raw = np.pad(raw, ((0, 0), (0, pad)))
        src = torch.from_numpy(raw).pin_memory().to(dev, non_blocking=True)
        dst = torch.empty((B, T, row_bytes + pad), dtype=torch.uint8, device=dev)
        bits = 0 if fill is None else int(np.float32(fill).view(np.uint32))
        with torch.cuda.device(dev):
            _lib.call("b200rl_pad_stack", _ptr(src), _ptr(off), _ptr(fill_rows), ctypes.c_uint32(bits), _ptr(dst),
                      B, T, (row_bytes + pad) // 4, _lib.stream_ptr())
        if pad:
            dst = dst[..., :row_bytes].contiguous()
        out[key] = dst.view(torch.from_numpy(np.empty(0, dt)).dtype).reshape((B, T) + tail)
    obs_keys = [k for k in out if isinstance(k, tuple)]
    if len(obs_keys) == 1 and obs_keys[0][1] is None:
        observation = out[obs_keys[0]]
    else:
        sample = episodes and _obs_leaves(_decode(episodes[0])[0]['observation'][_decode(episodes[0])[0]['turn'][0]])
        observation = {k[1]: out[k] for k in obs_keys}
        if isinstance(next(iter(sample)), int):
            observation = [observation[i] for i in sorted(observation)]
    return {
        'observation': observation,
        'policy': out['policy'], 'value': out['value'],
        'action': out['action'], 'outcome': oc_dev,
        'reward': out['reward'], 'return': out['return'],
        'episode_mask': out['episode_mask'],
        'turn_mask': out['turn_mask'], 'observation_mask': out['observation_mask'],
        'action_mask': out['action_mask'],
        'progress': out['progress'],
    }
