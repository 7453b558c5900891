"""GPU parity of the batched game engine (CUDA, through the C-ABI): trajectories,
legal lists/masks, observations (three views), rewards, terminals and outcomes must
be BIT-EXACT against (1) tapes recorded from the reference and (2) the C oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import games_c
from tests.tapeutil import legal_lists, tape_index, unpack

pytestmark = pytest.mark.gpu

ENVS = {'tictactoe': 'TicTacToe', 'parallel_tictactoe': 'ParallelTicTacToe', 'geister': 'Geister',
        'ci_geister': 'CIGeister'}


def replay_and_compare(env_key, t):
    from handyrl_b200.engine import BatchedEnv
    idx = tape_index(t)
    L, G = idx.shape
    geister = 'geister' in env_key
    nmain = 252 if geister else 27
    env = BatchedEnv(ENVS[env_key], G)
    A = env.A
    for k in range(L):
        rows = idx[k]
        alive = rows >= 0
        if k > 0:
            act = np.where(alive, t['action'][np.maximum(rows, 0)], -1).astype(np.int32)
            sel = np.where(alive, t['sel'][np.maximum(rows, 0)], 0).astype(np.int32)
            rew, term = env.step(act, sel, want_reward=True, want_terminal=True)
            rew = rew.cpu().numpy()
            if geister:
                np.testing.assert_array_equal(rew[alive, 0], t['reward'][rows[alive]])
                np.testing.assert_array_equal(rew[alive, 1], t['reward'][rows[alive]])
            np.testing.assert_array_equal(term.cpu().numpy()[alive], t['terminal'][rows[alive]])
        r = rows[alive]
        info = env.info().cpu().numpy()
        np.testing.assert_array_equal(info[alive, 1], t['terminal'][r])
        if env_key != 'parallel_tictactoe':
            np.testing.assert_array_equal(info[alive, 0], t['turn'][r])
        # legal list in reference order, mask as bits and as fp32 action_mask
        ids, cnt = env.legal_list()
        ids, cnt = ids.cpu().numpy(), cnt.cpu().numpy()
        ref_lists = legal_lists(t, r)
        got_lists = [list(ids[g][:cnt[g]]) for g in np.flatnonzero(alive)]
        assert got_lists == ref_lists
        ob = env.observe(view=-1, amask=True, bits=True)
        am = ob['action_mask'].cpu().numpy()[alive]
        bits = ob['legal_bits'].cpu().numpy().view(np.uint32)[alive]
        ref_am = np.full((len(r), A), 1e32, dtype=np.float32)
        for j, la in enumerate(ref_lists):
            ref_am[j, la] = 0
        np.testing.assert_array_equal(am, ref_am)
        got_bits = np.unpackbits(bits.view(np.uint8), axis=-1, bitorder='little')[:, :A]
        np.testing.assert_array_equal(got_bits, (ref_am == 0).astype(np.uint8))
        for key, view in (('N', -1), ('0', 0), ('1', 1)):
            o = ob if view == -1 else env.observe(view=view)
            np.testing.assert_array_equal(o['main'].cpu().numpy().reshape(G, -1)[alive], unpack(t['obs' + key][r], nmain))
            if geister:
                np.testing.assert_array_equal(o['scalar'].cpu().numpy()[alive], unpack(t['sc' + key][r], 18))
        # turn-player view == the explicit view of whoever is to move (not for ParallelTicTacToe)
        if env_key != 'parallel_tictactoe':
            o = env.observe(view=2)
            turn = t['turn'][r]
            ref = np.where(turn[:, None] == 0, unpack(t['obs0'][r], nmain), unpack(t['obs1'][r], nmain))
            np.testing.assert_array_equal(o['main'].cpu().numpy().reshape(G, -1)[alive], ref)
        # outcome once a game is finished
        done = alive & (np.r_[idx[k + 1:k + 2].reshape(-1, G)[0] if k + 1 < L else -np.ones(G)] < 0)
        if done.any():
            oc = env.outcome().cpu().numpy()
            np.testing.assert_array_equal(oc[done], t['outcome'][np.flatnonzero(done)])


@pytest.mark.parametrize("env_key", list(ENVS))
def test_reference_tapes(golden_dir, env_key):
    t = dict(np.load(os.path.join(golden_dir, "tape_%s.npz" % env_key)))   # NpzFile re-inflates on every access
    replay_and_compare(env_key, dict(t))


@pytest.mark.parametrize("env_key,n_games", [('tictactoe', 3000), ('parallel_tictactoe', 3000), ('geister', 1500),
                                             ('ci_geister', 300)])
def test_c_oracle_tapes(env_key, n_games):
    t = games_c.random_tapes(env_key, n_games, seed=11)
    replay_and_compare(env_key, t)


@pytest.mark.parametrize("env_key", list(ENVS))
def test_fused_step_observe_matches_separate_calls(env_key):
    from handyrl_b200.engine import BatchedEnv
    N = 1000
    a = BatchedEnv(ENVS[env_key], N)
    b = BatchedEnv(ENVS[env_key], N)
    gen = torch.Generator(device='cuda').manual_seed(1)
    out = a.step_observe(None, auto_reset=True)
    for k in range(40):
        ids, cnt = b.legal_list()
        pick = (torch.rand(N, device='cuda', generator=gen) * cnt).long().clamp(max=70)
        act = ids.gather(1, pick[:, None]).squeeze(1).int()
        sel = torch.randint(0, 2, (N,), device='cuda', generator=gen).int()
        out = a.step_observe(act, sel, auto_reset=True)
        rew, term = b.step(act, sel, want_reward=True, want_terminal=True)
        assert torch.equal(out['terminal'], term) and torch.equal(out['reward'], rew)
        assert torch.equal(out['outcome'], b.outcome())
        b.reset(term)
        ob = b.observe(view=2, amask=True)
        assert torch.equal(out['main'], ob['main'])
        assert torch.equal(out['action_mask'], ob['action_mask'])
        if ob['scalar'] is not None:
            assert torch.equal(out['scalar'], ob['scalar'])
        assert torch.equal(out['turn'], b.info()[:, 0])
        assert torch.equal(a.state[:N * a.state_bytes // 4], b.state[:N * b.state_bytes // 4])
