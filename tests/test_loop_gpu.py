"""GPU: the actor-learner loop of handyrl_b200/loop.py (BASELINE configs[2]/[3]; bench.py runs it as `extra.learner`) at
world size 1: generation on one stream beside training on another, the `minimum_episodes` gate (train.py:330-331), the
weight hand-off learner -> actor once per epoch (train.py:613 / worker.py:52) and the lr schedule per epoch
(train.py:396-398).  The NCCL collectives themselves are covered by the world-2 gloo test in tests/test_learner.py and by
bench.py under torchrun."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(**over):
    from handyrl_b200.loop import DEFAULT_ARGS
    return dict(copy.deepcopy(DEFAULT_ARGS), batch_size=64, forward_steps=8, **over)


def _flat(net):
    return torch.cat([p.detach().reshape(-1) for p in net.parameters()])


@pytest.mark.parametrize("env_name,overlap", [("TicTacToe", True), ("TicTacToe", False), ("ParallelTicTacToe", True),
                                               ("Geister", True)])
def test_actor_learner_epochs(env_name, overlap):
    from handyrl_b200.loop import ActorLearner
    geister = env_name == "Geister"
    n_games = 256 if geister else 1024
    a = _args(minimum_episodes=200, value_target="TD", policy_target="TD") if not geister else \
        _args(minimum_episodes=64, value_target="VTRACE", policy_target="UPGO")
    dev = torch.device("cuda", torch.cuda.current_device())
    al = ActorLearner(env_name, n_games, a, dev, overlap=overlap)
    assert torch.equal(_flat(al.actor), _flat(al.net))                 # the actor starts as a copy of the learner's net
    w0 = _flat(al.net).clone()

    # epoch 0: nothing finished yet -> the gate keeps every rank out of train() (and of its collectives)
    assert al.epoch(4, 2) is None
    torch.cuda.synchronize()
    assert torch.equal(_flat(al.net), w0)
    if geister:
        al.gen.rollout(220)                                            # epoch-0 policy fills the replay (games last ~160 plies)
        al.snap.copy_(al.gen.store.counters)

    lr0 = al.learner.optimizer.param_groups[0]["lr"]
    res = None
    for _ in range(6):
        res = al.epoch(8, 2) or res
    torch.cuda.synchronize()
    st = al.gen.stats()
    assert st["episodes"] >= a["minimum_episodes"]
    assert res is not None, "training never started although enough episodes finished"
    losses, dcnt = res
    assert dcnt > 0 and all(torch.isfinite(torch.tensor(v)) for v in losses.values())
    assert not torch.equal(_flat(al.net), w0)                          # the optimiser moved the weights ...
    assert torch.equal(_flat(al.actor), _flat(al.net))                 # ... and the actor got them at the epoch boundary
    assert al.learner.optimizer.param_groups[0]["lr"] != lr0           # schedule applied per epoch (train.py:396-398)
    assert al.learner.batch_cnt == 0                                   # ... and the epoch counters were reset
    # the trainer only ever saw complete episode rows: the snapshot trails the live counters of the running store
    assert int(al.snap[0]) <= st["episodes"]
    # recorded behaviour logits are usable: policy of live rows is finite where legal
    b = al.batch
    live = b["episode_mask"][..., 0, 0] > 0
    chosen = b["policy"].gather(-1, b["action"])[..., 0, 0]
    assert bool(live.any()) and bool((chosen[live] > -1e30).all())
