"""select_episode (handyrl/train.py:284-302): device kernel vs the Python restatement with the same
counter RNG (bit-exact), the host C entry point, and the whole device pipeline
rollout -> select_windows -> make_batch_windows against per-episode make_batch."""
import numpy as np
import pytest

from oracle import select_np

ARGS = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 16, 'compress_steps': 4,
        'maximum_episodes': 3000}


def _fake_table(n, E, rng, it_now):
    """finished-episode rows with increasing end_it, as the rollout kernels write them"""
    tab = np.zeros((E, 4), dtype=np.int32)
    end = np.sort(rng.integers(50, it_now, size=n))
    for k in range(n):
        steps = int(rng.integers(3, 203))
        steps = min(steps, int(end[k]) + 1)
        tab[k % E] = (rng.integers(0, 512), steps | (int(rng.integers(1, 4)) << 16), int(end[k]), 0)
    return tab


def test_python_restatement_is_deterministic_and_valid():
    rng = np.random.default_rng(1)
    tab = _fake_table(700, 1024, rng, 3000)
    w1 = select_np.select_windows(tab, 700, 500, 16, 99, 0, 3000, 4096, 64)
    w2 = select_np.select_windows(tab, 700, 500, 16, 99, 0, 3000, 4096, 64)
    assert np.array_equal(w1, w2)
    assert np.all(w1[:, 3] >= 1) and np.all(w1[:, 4] + w1[:, 3] <= w1[:, 5])
    assert not np.array_equal(w1, select_np.select_windows(tab, 700, 500, 16, 99, 4, 3000, 4096, 64))


@pytest.mark.gpu
@pytest.mark.parametrize("game,plies,D", [('Geister', 700, 1024), ('TicTacToe', 60, 64), ('Geister', 900, 256)])
def test_device_select_matches_restatement(game, plies, D):
    import torch
    from handyrl_b200.generation import BatchGenerator
    gen = BatchGenerator(game, 256, seed=5, depth=D, max_episodes=4096)
    gen.rollout(plies)
    a = dict(ARGS, maximum_episodes=1000)
    Bw = 512
    w = gen.select_windows(Bw, a, seed=77).cpu().numpy()
    tab = gen.store.ep_table.cpu().numpy()
    cur = int(gen.store.counters[5].item())
    assert cur > 0
    ref = select_np.select_windows(tab, cur, 1000, 16, 77 ^ 0x5E1EC7, 0, gen.store.it, D, Bw)
    assert np.array_equal(w, ref)
    # a second call advances the draw counter
    w2 = gen.select_windows(Bw, a, seed=77).cpu().numpy()
    ref2 = select_np.select_windows(tab, cur, 1000, 16, 77 ^ 0x5E1EC7, 256, gen.store.it, D, Bw)
    assert np.array_equal(w2, ref2) and not np.array_equal(w, w2)
    # every selected window lies inside the ring
    first = w[:, 1].astype(np.int64)
    assert np.all(gen.store.it - (first - w[:, 4]) <= D)


@pytest.mark.gpu
def test_select_before_any_episode_finished_gives_padded_windows():
    import torch
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.train import make_batch_windows
    gen = BatchGenerator('Geister', 64, seed=1, depth=64)
    gen.rollout(4)
    w = gen.select_windows(8, ARGS)
    assert np.array_equal(w.cpu().numpy(), np.tile(np.array([0, 0, 0, 0, 0, 1, 3, 0], np.int32), (8, 1)))
    b = make_batch_windows(gen.store, w, ARGS)
    assert float(b['episode_mask'].sum()) == 0 and float(b['progress'].min()) == 1.0
    assert float(b['action_mask'].min()) == float(np.float32(1e32))


@pytest.mark.gpu
@pytest.mark.parametrize("game", ['Geister', 'TicTacToe'])
def test_device_pipeline_equals_per_episode_make_batch(game):
    """windows drawn on the device and batched in one launch == the same windows fed to make_batch as
    episode dicts (which tests/test_make_batch_gpu.py pins to the reference)"""
    import torch
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.train import make_batch, make_batch_windows
    gen = BatchGenerator(game, 128, seed=3, depth=1024, max_episodes=8192)
    gen.rollout(500 if game == 'Geister' else 40)
    eps = gen.episodes()
    assert len(eps) > 20
    by_key = {(e['moment'].slot, e['moment'].first_it): e for e in eps}
    w = gen.select_windows(96, ARGS)
    got = make_batch_windows(gen.store, w, ARGS)
    wins = []
    for row in w.cpu().numpy():
        first = int(np.uint32(row[1])) | (int(np.uint32(row[2])) << 32)
        ep = by_key[(int(row[0]), first - int(row[4]))]
        assert ep['steps'] == int(row[5])
        wins.append({'args': {}, 'outcome': ep['outcome'], 'moment': ep['moment'], 'base': 0, 'start': int(row[4]),
                     'end': int(row[4] + row[3]), 'total': int(row[5])})
    ref = make_batch(wins, ARGS)

    def cmp(a, b, name):
        if isinstance(a, dict):
            for k in a:
                cmp(a[k], b[k], name + '.' + k)
        else:
            assert torch.equal(a, b), name
    for k in ref:
        cmp(got[k], ref[k], k)


def test_host_entry_point_matches_restatement():
    """b200rl_select_windows_host is a HOST function of the C-ABI library: same rule and RNG as the device kernel"""
    import ctypes
    from handyrl_b200 import _lib
    rng = np.random.default_rng(3)
    for (n, E, cur, max_eps, T, D, B, thr) in [(700, 1024, 700, 500, 16, 4096, 300, 1), (1024, 1024, 5000, 100000, 32, 600, 9000, 4),
                                               (5, 64, 5, 10, 8, 4096, 64, 2)]:
        tab = _fake_table(n, E, rng, 3000)
        out = np.zeros((B, 8), dtype=np.int32)
        _lib.call("b200rl_select_windows_host", ctypes.c_void_p(tab.ctypes.data), E, cur, max_eps, T, 99, 8, 3000, D,
                  ctypes.c_void_p(out.ctypes.data), B, thr)
        ref = select_np.select_windows(tab, cur, max_eps, T, 99, 8, 3000, D, B)
        assert np.array_equal(out, ref)


@pytest.mark.gpu
def test_host_replay_index_equals_device_selection():
    import torch
    from handyrl_b200.generation import BatchGenerator, HostReplayIndex
    gen = BatchGenerator('Geister', 512, seed=5, depth=1024, max_episodes=4096)
    gen.rollout(300)
    idx = HostReplayIndex(gen, max_new_rows=64, n_threads=2)      # small slab: a fetch brings at most 64 new rows
    c0 = idx.cursor
    gen.rollout(300)
    total = int(gen.store.counters[5].item())
    assert total - c0 > 3 * 64
    t1 = idx.fetch_async()
    t2 = idx.fetch_async()                                        # two read-backs in flight, same start: overlap
    assert idx.commit(t1) == c0 + 64 and idx.commit(t2) == c0 + 64
    for _ in range(100):
        if idx.cursor == total:
            break
        idx.commit(idx.fetch_async())                             # the rows past a slab come with the next fetches
    assert idx.cursor == total > 0
    assert torch.equal(idx.table[:min(idx.cursor, gen.store.E)], gen.store.ep_table.cpu()[:min(idx.cursor, gen.store.E)])
    a = dict(ARGS, maximum_episodes=2000)
    pin = torch.empty(256, 8, dtype=torch.int32).pin_memory()
    host = idx.select(256, a, pin, seed=7).clone()
    gen._draws = 0
    dev = gen.select_windows(256, a, seed=7).cpu()
    assert torch.equal(host, dev)


@pytest.mark.gpu
def test_lookahead_keeps_windows_inside_the_ring():
    """A pipelined caller draws windows, generates `lookahead` more plies, then builds the batch: with
    select(lookahead=K) every chosen episode still has its first ply in the trajectory ring after those K plies
    (without it, episodes that began within K plies of the ring's far edge could be chosen)."""
    import torch
    from handyrl_b200.generation import BatchGenerator, HostReplayIndex
    D, K = 256, 64
    gen = BatchGenerator('TicTacToe', 2048, seed=3, depth=D, max_episodes=1 << 16)
    gen.rollout(D + 40)                                   # the ring has wrapped: old episodes are gone
    idx = HostReplayIndex(gen, n_threads=2)
    a = dict(ARGS, maximum_episodes=60000, forward_steps=4)   # ~220 plies of finished episodes: some near the edge
    pin = torch.empty(4096, 8, dtype=torch.int32).pin_memory()
    w = idx.select(4096, a, pin, seed=1, lookahead=K).clone()
    it_use = gen.store.it + K                             # iteration counter when the batch would be built
    first = w[:, 1].to(torch.int64) & 0xFFFFFFFF | (w[:, 2].to(torch.int64) << 32)
    start_of_episode = first - w[:, 4].to(torch.int64)    # window start minus its offset in the episode
    assert bool((it_use - start_of_episode <= D).all())
    w0 = idx.select(4096, a, pin, seed=1, lookahead=0)
    first0 = w0[:, 1].to(torch.int64) & 0xFFFFFFFF | (w0[:, 2].to(torch.int64) << 32)
    assert bool((gen.store.it - (first0 - w0[:, 4].to(torch.int64)) <= D).all())
    # the case is not vacuous: without the lookahead some of the chosen episodes would be overwritten by then
    assert not bool((it_use - (first0 - w0[:, 4].to(torch.int64)) <= D).all())
