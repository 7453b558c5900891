"""GPU parity of make_batch: the device path (trajectory ring -> b200rl_make_batch) against
(1) golden outputs of the reference's make_batch on the same episodes and windows, and
(2) the numpy oracle on episodes exported from a device rollout.  Bar: bit-exact."""
import bz2
import os
import pickle
import random

import numpy as np
import pytest
import torch

from oracle import make_batch_np

pytestmark = pytest.mark.gpu
ENVS = {'tictactoe': 'TicTacToe', 'parallel_tictactoe': 'ParallelTicTacToe', 'geister': 'Geister',
        'ci_geister': 'CIGeister'}
CASES = ['ttt_turn', 'ttt_long', 'ttt_all', 'ttt_solo', 'pttt_solo', 'pttt_turn', 'geister_turn', 'geister_all',
         'cigeister_turn', 'geister_solo']


@pytest.fixture(scope="module")
def cases(golden_dir):
    with open(os.path.join(golden_dir, "batch_cases.pkl"), "rb") as f:
        return pickle.load(f)


def _flatten(batch):
    out = {}
    for k, v in batch.items():
        if isinstance(v, dict):
            for kk, vv in v.items():
                out['observation.' + kk] = vv.cpu().numpy()
        else:
            out[k] = v.cpu().numpy()
    return out


def replay_tapes(env_key, tapes):
    """slot g replays episode g; returns the generator and device episodes.  Tapes that carry the net outputs
    of every ply ('values' (steps,2), 'logits' (steps,PL,A)) fill the value / logit rings as well."""
    from handyrl_b200.generation import BatchGenerator, DeviceMoments
    G = len(tapes)
    L = max(t['steps'] for t in tapes)
    net = 'values' in tapes[0]
    gen = BatchGenerator(ENVS[env_key], G, depth=L + 2, record_values=net, record_logits=net)
    for k in range(L):
        act = np.array([t['actions'][k] if k < t['steps'] else -1 for t in tapes], np.int32)
        sel = np.array([t['sel'][k] if k < t['steps'] else 0 for t in tapes], np.int32)
        both = np.array([t['both'][k] if k < t['steps'] else [0, 0] for t in tapes], np.int32)
        values = logits = None
        if net:
            values = np.stack([t['values'][k] if k < t['steps'] else np.zeros(2, np.float32) for t in tapes])
            z = np.zeros_like(tapes[0]['logits'][0])
            logits = np.stack([t['logits'][k] if k < t['steps'] else z for t in tapes])
        gen.step(act, sel, both, values=values, logits=logits, auto_reset=False)
    eps = [{'args': {}, 'steps': t['steps'], 'outcome': {0: t['outcome'][0], 1: t['outcome'][1]},
            'moment': DeviceMoments(gen.store, g, 0, t['steps'])} for g, t in enumerate(tapes)]
    return gen, eps


@pytest.mark.parametrize("name", CASES)
def test_device_make_batch_matches_reference(cases, name):
    from handyrl_b200.train import make_batch
    c = cases[name]
    gen, eps = replay_tapes(c['env'], c['tapes'])
    # the episode table agrees with the tapes
    got = sorted((e['moment'].slot, e['steps'], e['outcome'][0]) for e in gen.episodes())
    assert got == sorted((g, t['steps'], t['outcome'][0]) for g, t in enumerate(c['tapes']))
    wins = [{'args': {}, 'outcome': eps[i]['outcome'], 'moment': eps[i]['moment'], 'base': 0, 'start': st,
             'end': ed, 'total': tot} for i, st, ed, tot in c['windows']]
    random.seed(c['seed'])
    out = _flatten(make_batch(wins, c['args']))
    assert set(out) == set(c['out'])
    for k, ref in c['out'].items():
        assert out[k].shape == ref.shape and out[k].dtype == ref.dtype, (k, out[k].shape, ref.shape, out[k].dtype)
        np.testing.assert_array_equal(out[k], ref, err_msg=k)


NET_CASES = ['ttt_net_turn', 'ttt_net_all', 'pttt_net_solo', 'geister_net_turn', 'geister_net_all']


@pytest.fixture(scope="module")
def net_cases(golden_dir):
    with open(os.path.join(golden_dir, "net_batch_cases.pkl.bz2"), "rb") as f:
        return pickle.loads(bz2.decompress(f.read()))


@pytest.mark.parametrize("name", NET_CASES)
def test_device_make_batch_with_recorded_net_outputs_matches_reference(net_cases, name):
    """A12 with a network in the loop (generation.py:36-57 -> train.py:60-82): the value ring read and
    k_policy_from_logits against the reference's Generator + make_batch run with a seeded reference net."""
    from handyrl_b200.train import make_batch
    c = net_cases[name]
    gen, eps = replay_tapes(c['env'], c['tapes'])
    wins = [{'args': {}, 'outcome': eps[i]['outcome'], 'moment': eps[i]['moment'], 'base': 0, 'start': st,
             'end': ed, 'total': tot} for i, st, ed, tot in c['windows']]
    random.seed(c['seed'])
    out = _flatten(make_batch(wins, c['args']))
    assert set(out) == set(c['out'])
    for k, ref in c['out'].items():
        assert out[k].shape == ref.shape and out[k].dtype == ref.dtype, (k, out[k].shape, ref.shape, out[k].dtype)
        np.testing.assert_array_equal(out[k], ref, err_msg=k)


@pytest.mark.parametrize("name", NET_CASES)
def test_reference_format_episodes_with_net_outputs_match_reference(net_cases, name):
    from handyrl_b200.train import make_batch
    c = net_cases[name]
    random.seed(c['seed'])
    out = _flatten(make_batch(c['ep_minimum'], c['args']))
    for k, ref in c['out'].items():
        np.testing.assert_array_equal(out[k], ref, err_msg=k)


@pytest.mark.parametrize("game,tbt,obs", [('Geister', True, False), ('Geister', True, True), ('TicTacToe', True, False),
                                          ('CIGeister', False, False)])
def test_rollout_export_roundtrip_against_oracle(game, tbt, obs):
    """device rollout -> export to the reference wire format -> numpy oracle make_batch == device make_batch"""
    from handyrl_b200.generation import BatchGenerator, export_episode
    from handyrl_b200.train import make_batch, select_episode
    gen = BatchGenerator(game, 64, seed=3, depth=512)
    gen.rollout(420 if 'Geister' in game else 40)
    eps = gen.episodes()
    assert len(eps) >= 32
    eps = eps[:24]
    args = {'turn_based_training': tbt, 'observation': obs, 'gamma': 0.8, 'forward_steps': 16, 'compress_steps': 4,
            'maximum_episodes': 1000}
    ref_eps = [export_episode(e, args) for e in eps]
    rng = random.Random(5)
    state = rng.getstate()
    dev_w = [select_episode(eps, args, rng) for _ in range(16)]
    rng.setstate(state)
    ref_w = [select_episode(ref_eps, args, rng) for _ in range(16)]
    random.seed(9)
    out = _flatten(make_batch(dev_w, args))
    random.seed(9)
    ref = make_batch_np.make_batch(ref_w, args)
    if obs:
        # The exported episodes were generated with observation=False: their non-turn moments are None, which the
        # reference's ALL mode pads (zero observation, omask 0), while the device re-encodes the non-turn player's
        # view from the record (omask 1, what observation=True generation yields - pinned by the golden *_all
        # cases).  On the turn player's rows, and for every per-ply scalar, the two must agree bit for bit.
        tm = ref['turn_mask'][..., 0] == 1                                  # (B, T, P)
        assert np.array_equal(out['turn_mask'], ref['turn_mask'])
        for k in ref:
            if k == 'observation_mask':
                live = ref['episode_mask'][:, :, 0, 0] == 1
                assert np.all(out[k][live] == 1) and np.all(out[k][~live] == 0)
            elif k.startswith('observation.') or k in ('policy', 'action_mask', 'action'):
                np.testing.assert_array_equal(out[k][tm], ref[k][tm], err_msg=k)
                if k.startswith('observation.'):                            # the reference's padding of None
                    assert not ref[k][~tm].any()
            else:
                np.testing.assert_array_equal(out[k], ref[k], err_msg=k)
        return
    for k in ref:
        np.testing.assert_array_equal(out[k], ref[k], err_msg=k)


def test_full_size_make_batch_properties():
    """BASELINE-size batch (B=4096 windows of T=16 from 65,536 live games): invariants without an oracle."""
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.train import make_batch, select_episode
    gen = BatchGenerator('Geister', 65536, seed=1, depth=256)
    gen.rollout(256)
    eps = gen.episodes()
    assert len(eps) > 20000
    eps = [e for e in eps if e['moment'].alive()]
    args = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 16,
            'compress_steps': 4, 'maximum_episodes': 100000}
    rng = random.Random(0)
    wins = [select_episode(eps, args, rng) for _ in range(4096)]
    b = make_batch(wins, args)
    em = b['episode_mask']
    assert b['policy'].shape == (4096, 16, 1, 214) and b['observation']['board'].shape == (4096, 16, 1, 7, 6, 6)
    live = em.bool().squeeze(-1).squeeze(-1)
    am = b['action_mask'][:, :, 0]
    assert bool(((am == 0) | (am == 1e32)).all())
    assert bool(((am == 0).sum(-1)[live] >= 1).all()) and bool(((am == 0).sum(-1)[~live] == 0).all())
    assert bool((b['policy'][:, :, 0][live] == -am[live]).all()) and bool((b['policy'][:, :, 0][~live] == 0).all())
    act = b['action'][:, :, 0, 0]
    assert bool((am.gather(-1, act[..., None]).squeeze(-1)[live] == 0).all())     # the played action was legal
    assert bool((b['turn_mask'].sum(2)[live] == 1).all())
    assert bool((b['observation']['board'][:, :, 0, 0][live] == 1).all())
    assert bool((b['observation']['board'][~live] == 0).all())
    assert bool((b['progress'][~live] == 1).all()) and bool((b['progress'][live] < 1).all())
    assert bool((b['value'][~live] == b['outcome'].expand(-1, 16, -1, -1)[~live]).all())


@pytest.mark.parametrize("name", CASES)
def test_reference_format_episodes_match_reference(cases, name):
    """the drop-in signature on the reference's own wire format (bz2-pickled moments)"""
    from handyrl_b200.train import make_batch
    c = cases[name]
    random.seed(c['seed'])
    out = _flatten(make_batch(c['ep_minimum'], c['args']))
    assert set(out) == set(c['out'])
    for k, ref in c['out'].items():
        assert out[k].shape == ref.shape and out[k].dtype == ref.dtype, (k, out[k].shape, ref.shape, out[k].dtype)
        np.testing.assert_array_equal(out[k], ref, err_msg=k)


def _custom_obs_episodes(n_ep, rng):
    """Reference-wire episodes of a made-up two-player game whose observation is a dict of a uint8 image with an odd
    row size (2x3 bytes), a float64 vector and a bool plane (the reference's to_torch keeps numpy dtypes,
    model.py:16-19)."""
    eps = []
    for e in range(n_ep):
        steps = int(rng.integers(2, 9))
        moments = []
        for t in range(steps):
            turn = t % 2
            m = {k: {0: None, 1: None} for k in ('observation', 'policy', 'action_mask', 'action', 'value', 'reward',
                                                'return')}
            m['observation'][turn] = {'img': rng.integers(0, 256, (2, 3), dtype=np.uint8),
                                      'vec': rng.standard_normal(5),
                                      'flag': rng.integers(0, 2, (3,)).astype(bool)}
            pol = rng.standard_normal(7).astype(np.float32)
            am = np.where(rng.random(7) < 0.5, 0, 1e32).astype(np.float32)
            m['policy'][turn], m['action_mask'][turn], m['action'][turn] = pol - am, am, int(rng.integers(0, 7))
            m['value'][turn] = np.array([rng.standard_normal()], np.float32)
            for p in (0, 1):
                m['reward'][p], m['return'][p] = float(rng.standard_normal()), float(rng.standard_normal())
            m['turn'] = [turn]
            moments.append(m)
        eps.append({'args': {}, 'steps': steps, 'outcome': {0: 1, 1: -1},
                    'moment': [bz2.compress(pickle.dumps(moments[i:i + 4])) for i in range(0, steps, 4)]})
    return eps


def test_reference_wire_path_keeps_any_observation_dtype():
    """pad_stack rows are raw bytes - uint8 / bool / float64 observation leaves must come back
    with their dtype, padded with zeros, bit-identical to the numpy oracle."""
    from handyrl_b200.train import make_batch, select_episode
    rng = np.random.default_rng(7)
    eps = _custom_obs_episodes(12, rng)
    args = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 6, 'compress_steps': 4,
            'maximum_episodes': 100}
    prng = random.Random(3)
    wins = [select_episode(eps, args, prng) for _ in range(20)]
    out = _flatten(make_batch(wins, args))
    ref = make_batch_np.make_batch(wins, args)
    assert out['observation.img'].dtype == np.uint8 and out['observation.vec'].dtype == np.float64
    assert out['observation.flag'].dtype == np.bool_
    assert set(out) == set(ref)
    for k in ref:
        assert out[k].shape == ref[k].shape and out[k].dtype == ref[k].dtype, (k, out[k].dtype, ref[k].dtype)
        np.testing.assert_array_equal(out[k], ref[k], err_msg=k)


@pytest.mark.parametrize("game,B,T,mode", [('Geister', 3, 5, 'turn'), ('Geister', 7, 9, 'all'), ('CIGeister', 33, 16, 'turn'),
                                           ('Geister', 5, 7, 'solo'), ('TicTacToe', 9, 3, 'turn'),
                                           ('ParallelTicTacToe', 5, 5, 'solo')])
def test_bulk_store_emit_equals_stg_emit(game, B, T, mode):
    """k_make_batch<G, TMA>: the staged bulk-store path (cp.async.bulk.global.shared::cta) must write exactly the
    bytes of the STG.128 path that the golden cases pin - including ragged last tiles with an odd number of rows
    (chunk sizes that are not multiples of 16 bytes end in plain stores)."""
    from handyrl_b200 import _lib
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.train import make_batch_windows
    gen = BatchGenerator(game, 96, seed=4, depth=512)
    gen.rollout(300 if 'Geister' in game else 40)
    args = {'turn_based_training': mode != 'solo', 'observation': mode == 'all', 'gamma': 0.8, 'forward_steps': T,
            'compress_steps': 4, 'maximum_episodes': 1000}
    w = gen.select_windows(B, args, seed=2)
    outs = []
    for knob in (0, 1):
        _lib.set_tuning('emit_tma', knob)
        try:
            outs.append(_flatten(make_batch_windows(gen.store, w, args)))
        finally:
            _lib.set_tuning('emit_tma', None)
    for k in outs[0]:
        np.testing.assert_array_equal(outs[0][k], outs[1][k], err_msg=k)
    assert float(np.abs(outs[0]['action_mask']).sum()) > 0
