"""compute_loss (handyrl/train.py:218-258) against reference losses recorded by tools/make_golden.py
(tests/golden/loss_cases.pkl), and the data-parallel plumbing (flat gradient all-reduce, weight broadcast)
on world_size-2 gloo."""
import os
import pickle
import socket

import numpy as np
import pytest
import torch


def _load(golden_dir):
    with open(os.path.join(golden_dir, "loss_cases.pkl"), "rb") as f:
        loss = pickle.load(f)
    with open(os.path.join(golden_dir, "batch_cases.pkl"), "rb") as f:
        batches = pickle.load(f)
    return loss, batches


def _net(key, weights):
    from handyrl_b200 import nets
    net = nets.GeisterNet() if key == 'geister' else nets.SimpleConv2dModel()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    return net


def test_golden_weights_load_into_our_nets(golden_dir):
    loss, _ = _load(golden_dir)
    for key, w in loss['weights'].items():
        net = _net(key, w)
        assert sum(p.numel() for p in net.parameters()) == (233832 if key == 'geister' else 29006)


CASES = ['ttt_turn:VTRACE:UPGO', 'ttt_all:TD:TD', 'ttt_long:MC:VTRACE', 'pttt_solo:UPGO:TD',
         'geister_turn:VTRACE:UPGO', 'geister_all:TD:VTRACE', 'geister_solo:UPGO:UPGO']


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("name", CASES)
def test_compute_loss_matches_reference(golden_dir, name, fused):
    from handyrl_b200.learner import compute_loss
    loss, batches = _load(golden_dir)
    c = loss['cases'][name]
    flat = batches[c['batch_case']]['out']
    dev = torch.device('cuda')
    batch = {k: torch.from_numpy(v).to(dev) for k, v in flat.items() if not k.startswith('observation.')}
    if any(k.startswith('observation.') for k in flat):
        batch['observation'] = {k.split('.', 1)[1]: torch.from_numpy(v).to(dev) for k, v in flat.items()
                                if k.startswith('observation.')}
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    net = _net(c['net'], loss['weights'][c['net']]).to(dev)
    net.train()
    B, P = batch['value'].shape[0], batch['value'].shape[2]
    hidden = None
    if hasattr(net, 'init_hidden'):
        hidden = net.init_hidden([B, P])
        hidden = tuple([t.to(dev) for t in part] for part in hidden)
    losses, dcnt = compute_loss(batch, net, hidden, c['args'], fused=fused)
    assert dcnt == c['dcnt']
    # fp32 tolerance: GPU convolutions and reductions order differently from the CPU reference
    for k, ref in c['losses'].items():
        assert float(losses[k].detach()) == pytest.approx(ref, rel=2e-3, abs=2e-3), k
    net.zero_grad()
    losses['total'].backward()
    gsq = sum(float((p.grad ** 2).sum()) for p in net.parameters() if p.grad is not None)
    assert gsq == pytest.approx(c['grad_sq_norm'], rel=2e-2)


# ---------------------------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _dp_worker(rank, world, port, q):
    import torch.distributed as dist
    from handyrl_b200 import nets
    from handyrl_b200.learner import all_ranks_have, allreduce_gradients, broadcast_weights
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    torch.manual_seed(100 + rank)                       # ranks start from different weights
    net = nets.SimpleConv2dModel()
    broadcast_weights(net, src=0)
    flat = torch.cat([p.detach().reshape(-1) for p in net.parameters()] + [b.reshape(-1).float() for b in net.buffers()])
    gathered = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    same_weights = all(torch.equal(gathered[0], g) for g in gathered)
    # rank-dependent gradients -> the mean on every rank
    x = torch.randn(4 + rank, 3, 3, 3)
    out = net(x)
    (out['policy'].sum() * (rank + 1) + out['value'].sum()).backward()
    local = [p.grad.clone() for p in net.parameters()]
    extra = allreduce_gradients(list(net.parameters()), extra=[100.0 * (rank + 1), 7.0])
    # the step's data count rides in the bucket and comes back as the mean over the ranks (same lr on every replica)
    ok = abs(extra[0] - 100.0 * (world + 1) / 2) < 1e-3 and abs(extra[1] - 7.0) < 1e-5
    for p, g in zip(net.parameters(), local):
        gs = [torch.zeros_like(g) for _ in range(world)]
        dist.all_gather(gs, g)
        ok = ok and torch.allclose(p.grad, sum(gs) / world, rtol=1e-6, atol=1e-7)
    # the minimum_episodes gate is collective: rank 1 has too few episodes -> nobody trains (loop.py)
    ok = ok and all_ranks_have(500, 400) and not all_ranks_have(torch.tensor([500 if rank == 0 else 10]), 400)
    q.put((rank, same_weights, ok))
    dist.destroy_process_group()


def test_allreduce_and_broadcast_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res), "weights differ after broadcast"
    assert all(r[2] for r in res), "gradients are not the mean over ranks"


def test_single_process_learner_helpers_are_noops():
    from handyrl_b200 import nets
    from handyrl_b200.learner import allreduce_gradients, broadcast_weights
    net = nets.SimpleConv2dModel()
    w = [p.detach().clone() for p in net.parameters()]
    broadcast_weights(net)
    allreduce_gradients(list(net.parameters()))
    assert allreduce_gradients(list(net.parameters()), extra=[3.0]) == [3.0]
    assert all(torch.equal(a, b) for a, b in zip(w, net.parameters()))
