"""Generate tests/golden/*.npz by running the REFERENCE itself (imported read-only
from /root/reference).  Runs only in the build container; the GPU box and the
tests read just the committed fixtures.

    PYTHONDONTWRITEBYTECODE=1 python tools/make_golden.py [targets|ttt|geister|batch|all]
"""
import os
import sys

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("HANDYRL_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

import numpy as np  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def gen_targets():
    import torch
    from handyrl.losses import compute_target  # the reference
    from tests.synth import target_inputs, TARGET_GOLDEN_CASES

    out = {}
    for ci, (B, T, P, head, rpp) in enumerate(TARGET_GOLDEN_CASES):
        d = target_inputs(B, T, P, head, rpp, seed=100 + ci)
        tt = {k: (torch.from_numpy(v) if isinstance(v, np.ndarray) else v) for k, v in d.items()}
        for k in ('values', 'returns', 'rewards', 'rhos', 'cs'):
            if d[k] is not None:
                out['c%d_%s' % (ci, k)] = d[k]
        for mode in ('MC', 'TD', 'UPGO', 'VTRACE'):
            tg, adv = compute_target(mode, tt['values'], tt['returns'], tt['rewards'], d['lmb'], d['gamma'],
                                     tt['rhos'], tt['cs'])
            out['c%d_%s_targets' % (ci, mode)] = tg.numpy()
            out['c%d_%s_advantages' % (ci, mode)] = adv.numpy()
    np.savez_compressed(os.path.join(GOLD, "targets.npz"), **out)
    print("targets.npz:", len(out), "arrays")


def _obs_to_bits(o):
    """reference observation -> (main bits packed uint8, scalar bits packed uint8)"""
    if isinstance(o, dict):
        main, scalar = np.ascontiguousarray(o['board']).reshape(-1), np.asarray(o['scalar']).reshape(-1)
    else:
        main, scalar = np.asarray(o).reshape(-1), np.zeros(0, np.float32)
    for a in (main, scalar):
        assert a.dtype == np.float32 and np.all((a == 0) | (a == 1))
    return (np.packbits(main.astype(np.uint8), bitorder='little'),
            np.packbits(scalar.astype(np.uint8), bitorder='little'))


def _play_tapes(env_module, n_games, parallel, seed0):
    """Random self-play through the REFERENCE env; records every state and action."""
    import random
    rec = dict(game=[], action=[], sel=[], turn=[], terminal=[], reward=[], legal_flat=[], legal_off=[0],
               obs0=[], obs1=[], obsN=[], sc0=[], sc1=[], scN=[], outcome=[], steps=[])
    for g in range(n_games):
        random.seed(seed0 + g)
        e = env_module.Environment()
        e.reset()

        def snap(action, sel, reward):
            rec['game'].append(g)
            rec['action'].append(action)
            rec['sel'].append(sel)
            t = e.turn()
            rec['turn'].append(t if isinstance(t, int) or hasattr(t, '__index__') else 0)
            rec['terminal'].append(1 if e.terminal() else 0)
            rec['reward'].append(reward)
            la = [int(a) for a in e.legal_actions(0)]
            rec['legal_flat'] += la
            rec['legal_off'].append(len(rec['legal_flat']))
            for key, pl in (('0', 0), ('1', 1), ('N', None)):
                m, sc = _obs_to_bits(e.observation(pl))
                rec['obs' + key].append(m)
                rec['sc' + key].append(sc)

        snap(-1, 0, 0.0)
        steps = 0
        while not e.terminal():
            actions = {p: random.choice(e.legal_actions(p)) for p in e.turns()}
            e.step(actions)
            steps += 1
            if parallel:
                color, act = e.record[-1]
                sel = 0 if color == e.BLACK else 1
            else:
                act, sel = list(actions.values())[0], 0
            r = e.reward()
            snap(int(act), sel, float(r.get(0, 0.0)) if r else 0.0)
        oc = e.outcome()
        rec['outcome'].append([oc[0], oc[1]])
        rec['steps'].append(steps)
    out = {}
    for k, v in rec.items():
        if k.startswith('obs') or k.startswith('sc'):
            out[k] = np.stack(v).astype(np.uint8)
        elif k in ('reward', 'outcome'):
            out[k] = np.asarray(v, dtype=np.float32)
        else:
            out[k] = np.asarray(v, dtype=np.int32)
    return out


def gen_games():
    import importlib
    for name, n_games, parallel in (('tictactoe', 200, False), ('parallel_tictactoe', 200, True),
                                    ('geister', 256, False), ('ci_geister', 64, False)):
        mod = importlib.import_module('handyrl.envs.' + name)
        out = _play_tapes(mod, n_games, parallel, seed0=1000)
        np.savez_compressed(os.path.join(GOLD, "tape_%s.npz" % name), **out)
        print("tape_%s.npz: %d games, %d states" % (name, n_games, len(out['game'])))


BATCH_CASES = [
    # name, env module, B, T, compress_steps, turn_based_training, observation, episodes
    ('ttt_turn', 'tictactoe', 12, 4, 4, True, False, 10),
    ('ttt_long', 'tictactoe', 6, 16, 4, True, False, 6),
    ('ttt_all', 'tictactoe', 8, 5, 2, True, True, 8),
    ('ttt_solo', 'tictactoe', 8, 6, 4, False, False, 8),
    ('pttt_solo', 'parallel_tictactoe', 10, 5, 4, False, False, 8),
    ('pttt_turn', 'parallel_tictactoe', 6, 6, 4, True, False, 6),
    ('geister_turn', 'geister', 8, 16, 4, True, False, 4),
    ('geister_all', 'geister', 4, 8, 4, True, True, 3),
    ('cigeister_turn', 'ci_geister', 4, 8, 4, True, False, 3),
    ('geister_solo', 'geister', 4, 8, 4, False, False, 3),
]


def gen_batch():
    import importlib
    import pickle
    import random
    from handyrl.generation import Generator
    from handyrl.model import ModelWrapper, RandomModel
    from handyrl.train import make_batch
    from handyrl_b200.train import select_episode

    cases = {}
    for ci, (name, envname, B, T, cs, tbt, obs, n_ep) in enumerate(BATCH_CASES):
        mod = importlib.import_module('handyrl.envs.' + envname)
        env = mod.Environment()
        args = {'turn_based_training': tbt, 'observation': obs, 'gamma': 0.8, 'forward_steps': T,
                'compress_steps': cs, 'maximum_episodes': 1000}
        gen = Generator(env, args)
        models = {p: ModelWrapper(RandomModel(env)) for p in env.players()}
        random.seed(500 + ci)
        episodes, tapes = [], []
        for _ in range(n_ep):
            ep = gen.generate(models, {})
            moments = sum([pickle.loads(__import__('bz2').decompress(ms)) for ms in ep['moment']], [])
            if envname == 'parallel_tictactoe':
                acts = [[m['action'][0], m['action'][1]] for m in moments]
                sel = [0 if color == env.BLACK else 1 for color, _ in env.record]
                landed = [a for _, a in env.record]
            else:
                landed = [m['action'][m['turn'][0]] for m in moments]
                acts = [[a, a] for a in landed]
                sel = [0] * len(landed)
            tapes.append({'actions': np.asarray(landed, np.int32), 'both': np.asarray(acts, np.int32),
                          'sel': np.asarray(sel, np.int32), 'steps': ep['steps'],
                          'outcome': [ep['outcome'][0], ep['outcome'][1]]})
            episodes.append(ep)
        rng = random.Random(900 + ci)
        picks = []
        for _ in range(B):
            w = select_episode(episodes, args, rng)
            idx = [i for i, e in enumerate(episodes) if e['outcome'] is w['outcome'] and e['steps'] == w['total']
                   and e['moment'][w['base'] // cs:] [:len(w['moment'])] == w['moment']][0]
            picks.append((idx, w))
        random.seed(77 + ci)
        out = make_batch([w for _, w in picks], args)
        random.seed(77 + ci)
        solo = [random.choice([0, 1]) for _ in picks] if not tbt else [0] * B
        flat = {}
        for k, v in out.items():
            if isinstance(v, dict):
                for kk, vv in v.items():
                    flat['observation.' + kk] = vv.numpy()
            else:
                flat[k] = v.numpy()
        cases[name] = {'env': envname, 'args': args, 'tapes': tapes,
                       'windows': [(idx, w['start'], w['end'], w['total']) for idx, w in picks],
                       'ep_minimum': [w for _, w in picks], 'solo': solo, 'seed': 77 + ci, 'out': flat}
        print(name, {k: v.shape for k, v in flat.items() if k in ('policy', 'value')})
    with open(os.path.join(GOLD, "batch_cases.pkl"), "wb") as f:
        pickle.dump(cases, f, protocol=4)
    print("batch_cases.pkl:", os.path.getsize(os.path.join(GOLD, "batch_cases.pkl")), "bytes")


NET_BATCH_CASES = [
    # name, env module, B, T, compress_steps, turn_based_training, observation, episodes
    ('ttt_net_turn', 'tictactoe', 10, 4, 4, True, False, 8),
    ('ttt_net_all', 'tictactoe', 8, 5, 2, True, True, 6),
    ('pttt_net_solo', 'parallel_tictactoe', 12, 5, 4, False, False, 8),
    ('geister_net_turn', 'geister', 6, 16, 4, True, False, 3),
    ('geister_net_all', 'geister', 4, 8, 4, True, True, 2),
]


def gen_netbatch():
    """Batches of episodes generated WITH A NETWORK in the loop (handyrl/generation.py:36-57 records the net's
    value and masked policy per moment; handyrl/train.py:60-82 copies them into the batch): the reference's
    Generator + make_batch with a seeded reference net.  The fixture keeps the action tape and the per-ply net
    outputs, so the device ring can be filled with exactly these numbers and b200rl_make_batch compared bit for bit
    on `policy` and `value` (and everything else)."""
    import bz2
    import importlib
    import pickle
    import random
    import torch
    from handyrl.generation import Generator
    from handyrl.model import ModelWrapper
    from handyrl.train import make_batch
    from handyrl_b200.train import select_episode

    cases = {}
    for ci, (name, envname, B, T, cs, tbt, obs, n_ep) in enumerate(NET_BATCH_CASES):
        mod = importlib.import_module('handyrl.envs.' + envname)
        env = mod.Environment()
        A = env.action_length()
        args = {'turn_based_training': tbt, 'observation': obs, 'gamma': 0.8, 'forward_steps': T,
                'compress_steps': cs, 'maximum_episodes': 1000}
        gen = Generator(env, args)
        torch.manual_seed(4000 + ci)
        net = env.net()()
        net.eval()
        models = {p: ModelWrapper(net) for p in env.players()}
        random.seed(600 + ci)
        episodes, tapes = [], []
        parallel = envname == 'parallel_tictactoe'
        for _ in range(n_ep):
            ep = gen.generate(models, {})
            moments = sum([pickle.loads(bz2.decompress(ms)) for ms in ep['moment']], [])
            PL = 2 if parallel else 1
            values = np.zeros((len(moments), 2), np.float32)
            logits = np.zeros((len(moments), PL, A), np.float32)
            for k, m in enumerate(moments):
                for p in (0, 1):
                    if m['value'][p] is not None:
                        values[k, p] = np.asarray(m['value'][p], np.float32).reshape(-1)[0]
                    if m['policy'][p] is not None:
                        legal = m['action_mask'][p] == 0
                        assert m['policy'][p].dtype == np.float32
                        logits[k, p if parallel else 0][legal] = m['policy'][p][legal]   # raw logits where legal
            if parallel:
                acts = [[m['action'][0], m['action'][1]] for m in moments]
                sel = [0 if color == env.BLACK else 1 for color, _ in env.record]
                landed = [a for _, a in env.record]
            else:
                landed = [m['action'][m['turn'][0]] for m in moments]
                acts = [[a, a] for a in landed]
                sel = [0] * len(landed)
            tapes.append({'actions': np.asarray(landed, np.int32), 'both': np.asarray(acts, np.int32),
                          'sel': np.asarray(sel, np.int32), 'steps': ep['steps'],
                          'outcome': [ep['outcome'][0], ep['outcome'][1]], 'values': values, 'logits': logits})
            episodes.append(ep)
        rng = random.Random(950 + ci)
        picks = []
        for _ in range(B):
            w = select_episode(episodes, args, rng)
            idx = [i for i, e in enumerate(episodes) if e['outcome'] is w['outcome'] and e['steps'] == w['total']
                   and e['moment'][w['base'] // cs:][:len(w['moment'])] == w['moment']][0]
            picks.append((idx, w))
        random.seed(88 + ci)
        out = make_batch([w for _, w in picks], args)
        flat = {}
        for k, v in out.items():
            if isinstance(v, dict):
                for kk, vv in v.items():
                    flat['observation.' + kk] = vv.numpy()
            else:
                flat[k] = v.numpy()
        cases[name] = {'env': envname, 'args': args, 'tapes': tapes,
                       'windows': [(idx, w['start'], w['end'], w['total']) for idx, w in picks],
                       'ep_minimum': [w for _, w in picks], 'seed': 88 + ci, 'out': flat}
        nz = float(np.abs(flat['value']).sum())
        print(name, {k: v.shape for k, v in flat.items() if k in ('policy', 'value')}, 'sum|value| %.3f' % nz)
        assert nz > 0
    blob = bz2.compress(pickle.dumps(cases, protocol=4))
    with open(os.path.join(GOLD, "net_batch_cases.pkl.bz2"), "wb") as f:
        f.write(blob)
    print("net_batch_cases.pkl.bz2:", len(blob), "bytes")


LOSS_CASES = [
    # batch case, value_target, policy_target
    ('ttt_turn', 'VTRACE', 'UPGO'), ('ttt_all', 'TD', 'TD'), ('ttt_long', 'MC', 'VTRACE'), ('pttt_solo', 'UPGO', 'TD'),
    ('geister_turn', 'VTRACE', 'UPGO'), ('geister_all', 'TD', 'VTRACE'), ('geister_solo', 'UPGO', 'UPGO'),
]


def gen_loss():
    """Reference compute_loss (handyrl/train.py:218-258) on the golden batches with seeded reference nets."""
    import importlib
    import pickle
    import torch
    from handyrl.model import ModelWrapper
    from handyrl.train import compute_loss
    with open(os.path.join(GOLD, "batch_cases.pkl"), "rb") as f:
        cases = pickle.load(f)
    out = {'weights': {}, 'cases': {}}
    for name, vt, pt in LOSS_CASES:
        c = cases[name]
        mod = importlib.import_module('handyrl.envs.' + c['env'])
        env = mod.Environment()
        net_key = 'geister' if 'geister' in c['env'] else 'ttt'
        torch.manual_seed(1234)
        net = env.net()()   # net() returns the class (train.py:436)
        if net_key not in out['weights']:
            out['weights'][net_key] = {k: v.numpy().copy() for k, v in net.state_dict().items()}
        model = ModelWrapper(net)
        model.train()
        args = dict(c['args'], **{'lambda': 0.7, 'value_target': vt, 'policy_target': pt,
                                  'entropy_regularization': 0.1, 'entropy_regularization_decay': 0.1})
        flat = c['out']
        batch = {k: torch.from_numpy(v) for k, v in flat.items() if not k.startswith('observation.')}
        if any(k.startswith('observation.') for k in flat):
            batch['observation'] = {k.split('.', 1)[1]: torch.from_numpy(v) for k, v in flat.items()
                                    if k.startswith('observation.')}
        B, P = batch['value'].shape[0], batch['value'].shape[2]
        hidden = model.init_hidden([B, P])
        losses, dcnt = compute_loss(batch, model, hidden, args)
        model.zero_grad()
        losses['total'].backward()
        gsq = sum(float((p.grad ** 2).sum()) for p in net.parameters() if p.grad is not None)
        out['cases'][name + ':' + vt + ':' + pt] = {
            'batch_case': name, 'net': net_key, 'args': args, 'dcnt': dcnt,
            'losses': {k: float(v) for k, v in losses.items()}, 'grad_sq_norm': gsq}
        print(name, vt, pt, {k: round(float(v), 4) for k, v in losses.items()}, dcnt, round(gsq, 3))
    with open(os.path.join(GOLD, "loss_cases.pkl"), "wb") as f:
        pickle.dump(out, f, protocol=4)
    print("loss_cases.pkl:", os.path.getsize(os.path.join(GOLD, "loss_cases.pkl")), "bytes")


SELECT_CASES = [
    # name, number of episodes, maximum_episodes, forward_steps, compress_steps, draws, seed
    ('short_buffer', 40, 1000, 16, 4, 1500, 11),          # accept rate ~ 1 everywhere
    ('steep', 300, 320, 8, 4, 1500, 12),                  # oldest episodes almost always rejected
    ('overfull', 260, 200, 16, 3, 1500, 13),              # len(episodes) > maximum_episodes (train.py:286-287)
    ('one_step_windows', 64, 100, 1, 1, 600, 14),
    ('long_window', 50, 60, 512, 8, 600, 15),             # forward_steps > every episode
]


def gen_select():
    """`Batcher.select_episode` of the reference (handyrl/train.py:284-302) called on a stub object, under
    random.seed(s): the draws (episode index, window, blocks) pin handyrl_b200.train.select_episode, and their
    histogram pins the law of the device / host-C window selection."""
    import pickle
    import random
    from types import SimpleNamespace
    from handyrl.train import Batcher
    from handyrl_b200.train import select_episode

    out = {}
    for name, n_ep, max_ep, T, cs, draws, seed in SELECT_CASES:
        rng = random.Random(seed)
        steps = [rng.randrange(3, 203) for _ in range(n_ep)]
        episodes = [{'args': {'id': i}, 'outcome': {0: 1, 1: -1}, 'steps': st,
                     'moment': [(i, b) for b in range((st + cs - 1) // cs)]} for i, st in enumerate(steps)]
        args = {'maximum_episodes': max_ep, 'forward_steps': T, 'compress_steps': cs}
        stub = SimpleNamespace(episodes=episodes, args=args)
        random.seed(seed)
        ref = [Batcher.select_episode(stub) for _ in range(draws)]
        mine_rng = random.Random(seed)
        rows = []
        for w in ref:
            m = select_episode(episodes, args, mine_rng)
            assert m == w, (name, m, w)
            rows.append((w['args']['id'], w['start'], w['end'], w['base'], w['total'], w['moment'][0][1],
                         len(w['moment'])))
        out[name] = {'steps': steps, 'args': args, 'seed': seed, 'rows': np.asarray(rows, np.int32)}
        print(name, 'ok', draws, 'draws; mean ep idx %.1f' % np.mean([r[0] for r in rows]))
    with open(os.path.join(GOLD, "select_cases.pkl"), "wb") as f:
        pickle.dump(out, f, protocol=4)


def gen_sampler():
    """The action draw of the reference's generation loop (handyrl/generation.py:48-53, util.py:61-63) on seeded
    states: (legal list in reference order, logits, the random() number `random.choices` consumed, action)."""
    import importlib
    import pickle
    import random
    from handyrl.util import softmax

    out = {}
    for envname, n_states in (('tictactoe', 700), ('geister', 1500)):
        mod = importlib.import_module('handyrl.envs.' + envname)
        env = mod.Environment()
        nprng = np.random.RandomState(31)
        rec = dict(tape_off=[0], tape=[], legal_off=[0], legal=[], logits=[], u=[], action=[])
        n = 0
        random.seed(4242)
        while n < n_states:
            env.reset()
            tape = []
            while not env.terminal() and n < n_states:
                player = env.turn()
                # sharp and flat logits alike, as float32 like a model output (model.py:50-53)
                p_ = (nprng.standard_normal(env.action_length()) * nprng.choice([0.1, 1.0, 4.0, 12.0])).astype(np.float32)
                legal_actions = env.legal_actions(player)
                action_mask = np.ones_like(p_) * 1e32
                action_mask[legal_actions] = 0
                p = p_ - action_mask
                state = random.getstate()
                action = random.choices(legal_actions, weights=softmax(p[legal_actions]))[0]
                after = random.getstate()
                random.setstate(state)
                u = random.random()                       # the one number random.choices drew
                assert random.getstate() == after
                rec['tape_off'].append(rec['tape_off'][-1])          # state = tape so far (filled below)
                rec['legal'] += [int(a) for a in legal_actions]
                rec['legal_off'].append(len(rec['legal']))
                rec['logits'].append(p_)
                rec['u'].append(u)
                rec['action'].append(int(action))
                tape.append(int(action))
                env.play(action, player)
                n += 1
            rec['tape'].append(tape)
        # states are addressed as (game index, ply): replaying tape[g][:ply] reaches state (g, ply)
        games, plies = [], []
        for g, tape in enumerate(rec['tape']):
            games += [g] * len(tape)
            plies += list(range(len(tape)))
        out[envname] = {
            'game': np.asarray(games, np.int32), 'ply': np.asarray(plies, np.int32),
            'tape_flat': np.asarray(sum(rec['tape'], []), np.int32),
            'tape_off': np.cumsum([0] + [len(t) for t in rec['tape']]).astype(np.int32),
            'legal': np.asarray(rec['legal'], np.int32), 'legal_off': np.asarray(rec['legal_off'], np.int32),
            'logits': np.stack(rec['logits']), 'u': np.asarray(rec['u'], np.float64),
            'action': np.asarray(rec['action'], np.int32)}
        print(envname, n, 'states,', len(rec['tape']), 'games')
    np.savez_compressed(os.path.join(GOLD, "sampler_cases.npz"),
                        **{e + '.' + k: v for e, d in out.items() for k, v in d.items()})


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    os.makedirs(GOLD, exist_ok=True)
    table = {k[4:]: v for k, v in list(globals().items()) if k.startswith("gen_")}
    for name, fn in table.items():
        if what in (name, "all"):
            fn()
