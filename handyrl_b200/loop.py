"""Actor-learner loop, one process per GPU (BASELINE.json configs 3 and 4).

Every rank owns a shard of the concurrent games (`game_id_base = rank * games`), generates with the current
network on its GPU (`BatchGenerator.play`), trains on windows of its own trajectory ring
(`select_windows` -> `make_batch_windows` -> `compute_loss`) and averages gradients over NCCL
(`DataParallelLearner`); weights and BatchNorm buffers are re-broadcast from rank 0 once per epoch — the
stand-in for the reference's model hand-off to the workers (train.py:613, worker.py:52).  This replaces the
roles of Learner/Trainer/Batcher/Worker for the hot path only; it has no sockets, queues or threads.

    python -m handyrl_b200.loop --env Geister --games 4096 --epochs 2
    torchrun --nproc-per-node 8 -m handyrl_b200.loop --env ParallelTicTacToe --value-target TD --policy-target TD
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist

from . import nets
from .generation import BatchGenerator
from .learner import DataParallelLearner, broadcast_weights
from .train import make_batch_windows

DEFAULT_ARGS = {  # reference config.yaml train_args
    'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 16, 'compress_steps': 4,
    'entropy_regularization': 1.0e-1, 'entropy_regularization_decay': 0.1, 'update_episodes': 200,
    'batch_size': 256, 'minimum_episodes': 400, 'maximum_episodes': 100000, 'lambda': 0.7,
    'policy_target': 'UPGO', 'value_target': 'VTRACE', 'seed': 0,
}


def make_net(env_name):
    return nets.GeisterNet() if 'Geister' in env_name else nets.SimpleConv2dModel()


class ActorLearner:
    def __init__(self, env_name, n_games, args, device, rank=0, world=1):
        self.args, self.device, self.rank, self.world = args, device, rank, world
        if env_name == 'ParallelTicTacToe':
            args['turn_based_training'] = False          # simultaneous game: one player per episode (train.py:57-58)
        torch.manual_seed(args['seed'])
        self.net = make_net(env_name).to(device)
        depth = 1024 if 'Geister' in env_name else 256
        self.gen = BatchGenerator(env_name, n_games, device=device, seed=args['seed'], game_id_base=rank * n_games,
                                  depth=depth, max_episodes=1 << 20, record_values=True, record_logits=True)
        self.learner = DataParallelLearner(self.net, args)
        self.n_games = n_games
        self.hidden = self.gen.init_hidden(self.net)      # per player: (n_games, 2, ...)
        self.batch = None
        self.simultaneous = env_name == 'ParallelTicTacToe'

    def _init_hidden(self, lead):
        if not hasattr(self.net, 'init_hidden'):
            return None
        h = self.net.init_hidden(lead)
        return tuple([t.to(self.device) for t in part] for part in h)

    @torch.no_grad()
    def generate(self, plies):
        self.net.eval()
        self.hidden = self.gen.play(self.net, plies, self.hidden)

    def train(self, steps):
        a = self.args
        stats = None
        for _ in range(steps):
            w = self.gen.select_windows(a['batch_size'], a)
            self.batch = make_batch_windows(self.gen.store, w, a, out=self.batch)
            B, P = self.batch['value'].shape[0], self.batch['value'].shape[2]
            stats = self.learner.train_step(self.batch, self._init_hidden([B, P]))
        return stats

    def epoch(self, plies, train_steps):
        self.generate(plies)
        out = None
        if self.gen.stats()['episodes'] >= min(self.args['minimum_episodes'], 1):
            out = self.train(train_steps)
        broadcast_weights(self.net, 0)
        return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--env', default='Geister')
    ap.add_argument('--games', type=int, default=4096)
    ap.add_argument('--epochs', type=int, default=2)
    ap.add_argument('--plies', type=int, default=64)
    ap.add_argument('--train-steps', type=int, default=4)
    ap.add_argument('--batch-size', type=int, default=256)
    ap.add_argument('--value-target', default='VTRACE')
    ap.add_argument('--policy-target', default='UPGO')
    a = ap.parse_args()
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    args = dict(DEFAULT_ARGS, batch_size=a.batch_size, value_target=a.value_target, policy_target=a.policy_target)
    al = ActorLearner(a.env, a.games, args, dev, rank, world)
    for ep in range(a.epochs):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = al.epoch(a.plies, a.train_steps)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        st = al.gen.stats()
        if rank == 0:
            print(json.dumps({'epoch': ep, 'seconds': round(dt, 3), 'env_steps_per_s_all_ranks':
                              round(a.games * a.plies * world / dt), 'episodes_rank0': st['episodes'],
                              'losses': res[0] if res else None, 'data_count': res[1] if res else 0,
                              'world': world}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
