"""GPU: the acting half of a self-play ply with a real policy (generation.py:48-61) — sampling rule against
the Python restatement for given uniforms, recording, and net-in-the-loop play feeding make_batch."""
import numpy as np
import pytest
import torch

from oracle import sample_np

pytestmark = pytest.mark.gpu
ARGS = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 8, 'compress_steps': 4,
        'maximum_episodes': 100000}


@pytest.mark.parametrize("game,warm", [('Geister', 60), ('Geister', 1), ('TicTacToe', 3), ('CIGeister', 120)])
def test_sampling_rule_matches_restatement(game, warm):
    from handyrl_b200.generation import BatchGenerator
    N = 1024
    gen = BatchGenerator(game, N, seed=11, depth=256)
    gen.rollout(warm)
    gen.observe_turn()                                   # restarts finished games, like act() will
    ids, cnt = gen.env.legal_list()
    ids, cnt = ids.cpu().numpy(), cnt.cpu().numpy()
    g = torch.Generator(device='cuda').manual_seed(5)
    logits = torch.randn(N, gen.env.A, device='cuda', generator=g) * 3
    u = torch.rand(N, dtype=torch.float64, device='cuda', generator=g)
    act, term = gen.act(logits, u=u)
    act, lg, uu = act.cpu().numpy()[:, 0], logits.cpu().numpy(), u.cpu().numpy()
    bad = 0
    for i in range(N):
        legal = ids[i, :cnt[i]].tolist()
        ref, cum = sample_np.choose(legal, lg[i], float(uu[i]))
        assert act[i] in legal
        if act[i] != ref:   # only possible when u*total sits within rounding of a cumulative-weight boundary
            x = uu[i] * cum[-1]
            assert min(abs(x - c) for c in cum) < 1e-6
            bad += 1
    assert bad <= 2


def test_parallel_tictactoe_samples_both_players():
    from handyrl_b200.generation import BatchGenerator
    N = 512
    gen = BatchGenerator('ParallelTicTacToe', N, seed=3, depth=64)
    gen.observe_turn()
    logits = torch.randn(N, 2, 9, device='cuda')
    u = torch.rand(N, 2, dtype=torch.float64, device='cuda')
    act, term = gen.act(logits, u=u)
    assert act.shape == (N, 2) and int(act.min()) >= 0 and int(act.max()) < 9
    rec = gen.store.traj[:, 0, 4:].cpu().numpy().view(np.uint32)   # words 4..7 of the records of iteration 0
    landed, sel = rec[:, 1] & 0xFFFF, (rec[:, 1] >> 16) & 1
    a = act.cpu().numpy()
    assert np.array_equal(landed, a[np.arange(N), sel])
    assert np.array_equal(rec[:, 3] & 0xFF, a[:, 0]) and np.array_equal((rec[:, 3] >> 8) & 0xFF, a[:, 1])
    assert 0.3 < sel.mean() < 0.7


@pytest.mark.parametrize("game", ['TicTacToe', 'Geister'])
def test_play_with_network_feeds_make_batch(game):
    from handyrl_b200 import nets
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.train import make_batch_windows
    N = 256
    torch.manual_seed(0)
    net = (nets.GeisterNet() if game == 'Geister' else nets.SimpleConv2dModel()).cuda().eval()
    gen = BatchGenerator(game, N, seed=9, depth=512, max_episodes=1 << 14, record_values=True, record_logits=True)
    assert gen.store.logits.shape == (N, 512, 1, 72 if game == 'Geister' else 12)
    hidden = gen.init_hidden(net)          # per player: leaves (N, 2, ...); None for the feed-forward net
    assert (hidden is None) == (game != 'Geister')
    plies = 260 if game == 'Geister' else 24
    gen.play(net, plies, hidden)
    st = gen.stats()
    assert st['env_steps'] == N * plies and st['episodes'] > 0
    assert st['episodes'] == st['wins0'] + st['wins1'] + st['draws']
    w = gen.select_windows(64, ARGS)
    b = make_batch_windows(gen.store, w, ARGS)
    live = b['episode_mask'][..., 0, 0] > 0                              # (B, T)
    assert bool(live.any())
    # recorded behaviour policy = logits - action_mask: the chosen action is always legal (finite logit)
    chosen = b['policy'].gather(-1, b['action'])[..., 0, 0]
    assert bool((chosen[live] > -1e30).all())
    assert bool((b['action_mask'].gather(-1, b['action'])[..., 0, 0][live] == 0).all())
    # values were recorded for the turn player only
    tm, v = b['turn_mask'][live], b['value'][live]
    assert bool((v[tm == 0] == 0).all()) and float(v.abs().max()) <= 1.0
    assert float(v[tm == 1].abs().sum()) > 0


def test_recorded_logits_are_what_the_learner_recomputes():
    """the behaviour logits/values recorded by `play` must come from the same per-player
    recurrent context that `forward_prediction` (train.py:160-175) rebuilds from the batch: windows that start at
    ply 0 of an episode, zero (B, P) hidden state, same weights -> the recomputed masked policy and the values
    equal the recorded ones (up to cuDNN's batch-shape dependent summation order)."""
    from handyrl_b200 import nets
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.learner import forward_prediction
    from handyrl_b200.train import make_batch_windows
    N, T = 96, 12
    torch.manual_seed(3)
    net = nets.GeisterNet().cuda().eval()
    gen = BatchGenerator('Geister', N, seed=21, depth=64, record_values=True, record_logits=True)
    gen.play(net, T, gen.init_hidden(net))
    # windows = plies 0..T-1 of every slot's first episode (no game ends within 12 plies)
    win = torch.zeros(N, 8, dtype=torch.int32, device='cuda')
    win[:, 0] = torch.arange(N)
    win[:, 3], win[:, 5], win[:, 6] = T, 200, 3
    args = dict(ARGS, forward_steps=T)
    b = make_batch_windows(gen.store, win, args)
    assert float(b['episode_mask'].min()) == 1.0
    h0 = tuple([t.cuda() for t in part] for part in net.init_hidden([N, 2]))
    with torch.no_grad():
        out = forward_prediction(net, h0, b, args)
    legal = b['action_mask'] == 0
    assert bool((b['policy'][~legal] == -1e32).all())
    d = (out['policy'] - b['policy'])[legal].abs().max().item()
    assert d < 2e-4, d
    tm = b['turn_mask'] > 0
    dv = (out['value'] - b['value'])[tm].abs().max().item()
    assert dv < 2e-4, dv
    # ... and it is NOT what a state shared by both players produces: at ply 2 (player 0's second observation, same
    # board in both runs) the turn player's own state differs from "whatever the opponent just wrote"
    gen2 = BatchGenerator('Geister', N, seed=21, depth=64, record_values=True, record_logits=True)
    shared = tuple([t[:, 0].clone() for t in part] for part in gen.init_hidden(net))
    for _ in range(T):
        o = gen2.observe_turn()
        r = net({'scalar': o['scalar'], 'board': o['main']}, shared)
        shared = r['hidden']
        gen2.act(r['policy'].float())
    b2 = make_batch_windows(gen2.store, win, args)
    same = (b2['policy'][:, :2] == b['policy'][:, :2]).all()
    assert bool(same)            # the first ply of each player starts from zeros either way
    assert bool((b2['action_mask'][:, 2] == b['action_mask'][:, 2]).all())
    assert (b2['policy'][:, 2] - b['policy'][:, 2])[legal[:, 2]].abs().max() > 1e-4


def test_parallel_tictactoe_records_both_players_logits():
    """SOLO windows of player 1 must carry player 1's behaviour logits."""
    from handyrl_b200.generation import BatchGenerator
    from handyrl_b200.train import make_batch_windows
    N, T = 128, 4
    gen = BatchGenerator('ParallelTicTacToe', N, seed=2, depth=16, record_values=True, record_logits=True)
    g = torch.Generator(device='cuda').manual_seed(1)
    fed = []
    for _ in range(T):
        gen.observe_turn()
        lg = torch.randn(N, 2, 9, device='cuda', generator=g)
        fed.append(lg)
        gen.act(lg, values=torch.rand(N, 2, device='cuda', generator=g))
    fed = torch.stack(fed, 1)                                         # (N, T, 2, 9)
    args = dict(ARGS, turn_based_training=False, forward_steps=T)
    for solo in (0, 1):
        win = torch.zeros(N, 8, dtype=torch.int32, device='cuda')
        win[:, 0] = torch.arange(N)
        win[:, 3], win[:, 5], win[:, 6], win[:, 7] = 3, 9, 3, solo    # 3 plies: no game can end before its 5th ply... of one colour
        b = make_batch_windows(gen.store, win, args)
        legal = b['action_mask'][:, :3, 0] == 0
        assert bool((b['policy'][:, :3, 0][legal] == fed[:, :3, solo][legal]).all())
        assert bool((b['policy'][:, :3, 0][~legal] == -1e32).all())
