"""Episode-to-batch assembly on the GPU — drop-in for `make_batch` of `handyrl/train.py`.

`make_batch(episodes, args)` keeps the reference signature and returns the same 12-key dict
with the same shapes and dtypes (handyrl/train.py:33-133), as CUDA tensors on the device that
holds the trajectories (so the reference's `to_gpu(batch)` at train.py:377 is a no-op move).

Two input forms are accepted:
  * device episodes from `handyrl_b200.generation.BatchGenerator` (`moment` is a DeviceMoments
    handle): windows are re-encoded straight from the 32-byte trajectory records by one kernel
    (`b200rl_make_batch`); nothing is decompressed or copied through the host;
  * reference episodes (`moment` is a list of bz2-pickled blocks, generation.py:82-85): the
    blocks are decoded on the host (that decode cannot move to the GPU), the per-episode
    arrays are uploaded ragged and padded/stacked by `b200rl_pad_stack` on the device.
"""
import bz2
import ctypes
import pickle
import random

import numpy as np
import torch

from . import _lib
from .engine import _ptr
from .generation import DeviceMoments

_RET_CACHE = {}
_RET_DEV_CACHE = {}


def _return_table(gamma, reward, n):
    """ret[m] = discounted return with m plies left, accumulated in double like
    generation.py:73-77 (`ret = (reward or 0) + gamma * ret`), rounded to fp32 at the end."""
    key = (float(gamma), reward, n)
    if key not in _RET_CACHE:
        tab, ret = [0.0], 0.0
        for _ in range(n):
            ret = (reward or 0) + gamma * ret
            tab.append(ret)
        _RET_CACHE[key] = np.asarray(tab, dtype=np.float32)
    return _RET_CACHE[key]


def _return_table_dev(gamma, reward, n, dev):
    key = (float(gamma), reward, n, str(dev))
    if key not in _RET_DEV_CACHE:
        _RET_DEV_CACHE[key] = torch.from_numpy(_return_table(gamma, reward, n)).to(dev)
    return _RET_DEV_CACHE[key]


def select_episode(episodes, args, rng=random):
    """Window selection of `Batcher.select_episode` (handyrl/train.py:284-302), same draws in the
    same order from `rng` (the reference uses the global `random` module)."""
    while True:
        ep_idx = rng.randrange(min(len(episodes), args['maximum_episodes']))
        accept_rate = 1 - (len(episodes) - 1 - ep_idx) / args['maximum_episodes']
        if rng.random() < accept_rate:
            break
    ep = episodes[ep_idx]
    turn_candidates = 1 + max(0, ep['steps'] - args['forward_steps'])
    st = rng.randrange(turn_candidates)
    ed = min(st + args['forward_steps'], ep['steps'])
    st_block = st // args['compress_steps']
    ed_block = (ed - 1) // args['compress_steps'] + 1
    ep_minimum = {
        'args': ep['args'], 'outcome': ep['outcome'],
        'moment': ep['moment'][st_block:ed_block],
        'base': st_block * args['compress_steps'],
        'start': st, 'end': ed, 'total': ep['steps']
    }
    return ep_minimum


def _win_code(outcome):
    o0 = outcome[0]
    return 1 if o0 > 0 else (2 if o0 < 0 else 3)


def _make_batch_device(episodes, args):
    store = episodes[0]['moment'].store
    env = store.env
    dev = env.device
    B = len(episodes)
    mode = batch_mode(args)
    win = np.zeros((B, 8), dtype=np.int64)
    for b, ep in enumerate(episodes):
        mo = ep['moment']
        if mo.store is not store:
            raise _lib.B200RLError("make_batch: all device episodes of a batch must live in one TrajectoryStore")
        if not mo.alive():
            raise _lib.B200RLError("make_batch: episode of slot %d was overwritten in the trajectory ring "
                                   "(raise TrajectoryStore depth)" % mo.slot)
        first = mo.first_it + ep['start']
        win[b] = (mo.slot, first & 0xFFFFFFFF, first >> 32, ep['end'] - ep['start'], ep['start'], ep['total'],
                  _win_code(ep['outcome']),
                  random.choice([0, 1]) if mode == 2 else 0)   # train.py:57-58
    wt = torch.from_numpy((win & 0xFFFFFFFF).astype(np.uint32).view(np.int32)).to(dev)
    return make_batch_windows(store, wt, args)


def batch_mode(args):
    """TURN (train.py:64-68), ALL (train.py:69-73) or SOLO (train.py:57-58)"""
    if args['turn_based_training'] and not args['observation']:
        return 0
    return 1 if args['turn_based_training'] else 2


def make_batch_windows(store, windows, args, out=None):
    """make_batch over device windows: `windows` is the (B,8) int32 CUDA tensor of
    `b200rl_select_windows` / `_make_batch_device` (slot, first_it lo/hi, len, start, total, winner, solo).
    `out` may carry the tensors of a previous call of the same shape to be overwritten in place."""
    env = store.env
    dev = env.device
    B, T = int(windows.shape[0]), int(args['forward_steps'])
    mode = batch_mode(args)
    Po = 2 if mode == 1 else 1
    Pv = 1 if mode == 2 else 2
    ret_tab = _return_table_dev(args['gamma'], -0.01 if env.has_reward else None, env.max_steps + 1, dev)
    if out is None:
        f32 = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        obs_main = f32(B, T, Po, *env.obs_shape)
        obs_scalar = f32(B, T, Po, *env.scalar_shape) if env.scalar_shape else None
        out = {
            'observation': {'scalar': obs_scalar, 'board': obs_main} if obs_scalar is not None else obs_main,
            'policy': f32(B, T, Po, env.A), 'value': f32(B, T, Pv, 1),
            'action': torch.empty(B, T, Po, 1, dtype=torch.int64, device=dev), 'outcome': f32(B, 1, Pv, 1),
            'reward': f32(B, T, Pv, 1), 'return': f32(B, T, Pv, 1),
            'episode_mask': f32(B, T, 1, 1),
            'turn_mask': f32(B, T, Pv, 1), 'observation_mask': f32(B, T, Pv, 1),
            'action_mask': f32(B, T, Po, env.A),
            'progress': f32(B, T, 1),
        }
    obs = out['observation']
    obs_main, obs_scalar = (obs['board'], obs['scalar']) if isinstance(obs, dict) else (obs, None)
    with torch.cuda.device(dev):
        _lib.call("b200rl_make_batch", env.game, _ptr(store.traj), store.D, store.N, _ptr(store.values),
                  _ptr(store.logits), _ptr(windows), B, T, mode, 1 if args['observation'] else 0, _ptr(ret_tab),
                  ret_tab.numel(), _ptr(obs_main), _ptr(obs_scalar), _ptr(out['policy']), _ptr(out['action_mask']),
                  _ptr(out['action']), _ptr(out['value']), _ptr(out['reward']), _ptr(out['return']),
                  _ptr(out['outcome']), _ptr(out['episode_mask']), _ptr(out['turn_mask']),
                  _ptr(out['observation_mask']), _ptr(out['progress']), _lib.stream_ptr())
    return out


def make_batch(episodes, args):
    """Making training batch

    Args:
        episodes (Iterable): list of episodes
        args (dict): training configuration

    Returns:
        dict: PyTorch input and target tensors, basic data shape (B, T, P, ...)
    """
    episodes = list(episodes)
    if len(episodes) == 0:
        raise _lib.B200RLError("make_batch: empty episode list")
    if all(isinstance(ep['moment'], DeviceMoments) for ep in episodes):
        return _make_batch_device(episodes, args)
    from .train_host import make_batch_from_reference_episodes
    return make_batch_from_reference_episodes(episodes, args)
