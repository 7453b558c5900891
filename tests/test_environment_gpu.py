"""GPU: the single-game `Environment` facades (handyrl_b200.envs.*), tested the way the reference
tests its environments (tests/test_environment.py there: property / local loop / network mirror
loop with the legal-action set assertion), plus value parity against reference tapes."""
import importlib
import os
import random

import numpy as np
import pytest

from tests.tapeutil import unpack

pytestmark = pytest.mark.gpu
ENVS = ['tictactoe', 'geister', 'parallel_tictactoe', 'ci_geister']


def _mod(env):
    return importlib.import_module('handyrl_b200.envs.' + env)


@pytest.mark.parametrize('env', ENVS)
def test_environment_property(env):
    e = _mod(env).Environment()
    assert e.players() == [0, 1]
    assert e.action_length() == (214 if 'geister' in env else 9)
    assert isinstance(str(e), str)
    assert isinstance(e.net()(), __import__('torch').nn.Module)


@pytest.mark.parametrize('env', ENVS)
def test_environment_local(env):
    random.seed(1)
    e = _mod(env).Environment()
    for _ in range(6):
        e.reset()
        while not e.terminal():
            actions = {}
            for player in e.turns():
                actions[player] = random.choice(e.legal_actions(player))
            assert not e.step(actions)
            e.reward()
        assert set(e.outcome().keys()) == {0, 1}


@pytest.mark.parametrize('env', ENVS)
def test_environment_network(env):
    """server env + one mirror per player kept in sync through diff_info/update"""
    random.seed(2)
    mod = _mod(env)
    e = mod.Environment()
    es = {p: mod.Environment() for p in e.players()}
    for _ in range(4):
        e.reset()
        for p, e_ in es.items():
            e_.update(e.diff_info(p), True)
        while not e.terminal():
            actions = {}
            for player in e.turns():
                assert set(e.legal_actions(player)) == set(es[player].legal_actions(player))
                action = random.choice(es[player].legal_actions(player))
                actions[player] = es[player].action2str(action, player)
            actions = {p: e.str2action(a, p) for p, a in actions.items()}
            e.step(actions)
            for p, e_ in es.items():
                e_.update(e.diff_info(p), False)
            e.reward()
        e.outcome()


@pytest.mark.parametrize('env', ENVS)
def test_facade_replays_reference_tape(golden_dir, env):
    t = dict(np.load(os.path.join(golden_dir, "tape_%s.npz" % env)))   # NpzFile re-inflates on every access
    geister = 'geister' in env
    nmain = 252 if geister else 27
    e = _mod(env).Environment()
    n_games = 3
    game = -1
    for i in range(len(t['game'])):
        if t['game'][i] >= n_games:
            break
        if t['game'][i] != game:
            game = t['game'][i]
            e.reset()
        elif env == 'parallel_tictactoe':
            e._step(int(t['action'][i]), int(t['sel'][i]))
        else:
            e.play(int(t['action'][i]), None)
        assert e.terminal() == bool(t['terminal'][i])
        legal = [int(a) for a in t['legal_flat'][t['legal_off'][i]:t['legal_off'][i + 1]]]
        assert e.legal_actions(0) == legal
        for key, pl in (('0', 0), ('1', 1), ('N', None)):
            o = e.observation(pl)
            main = o['board'] if geister else o
            assert main.dtype == np.float32 and main.shape == ((7, 6, 6) if geister else (3, 3, 3))
            np.testing.assert_array_equal(main.reshape(-1), unpack(t['obs' + key][i], nmain))
            if geister:
                np.testing.assert_array_equal(o['scalar'], unpack(t['sc' + key][i], 18))
        if i + 1 == len(t['game']) or t['game'][i + 1] != game:
            oc = e.outcome()
            assert [oc[0], oc[1]] == list(t['outcome'][game])
            if geister:
                assert e.reward() == {0: -0.01, 1: -0.01}


def test_geister_string_codec_roundtrip():
    e = _mod('geister').Environment()
    e.play(144 + 5)
    e.play(144 + 60)
    for _ in range(30):
        p = e.turn()
        for a in e.legal_actions(p):
            assert e.str2action(e.action2str(a, p), p) == a
        e.play(random.choice(e.legal_actions(p)))
        if e.terminal():
            break
    assert 'record = ' in str(e) and 'color = ' in str(e)


def test_make_env_by_name_and_path():
    from handyrl_b200.environment import make_env, prepare_env
    prepare_env({'env': 'TicTacToe'})
    assert make_env({'env': 'TicTacToe'}).action_length() == 9
    assert make_env({'env': 'handyrl_b200.envs.geister'}).action_length() == 214
    assert make_env({'env': 'CIGeister', 'id': 3}).action_length() == 214
