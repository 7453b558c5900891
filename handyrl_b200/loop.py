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
from .learner import DataParallelLearner, all_ranks_have, broadcast_weights
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
    """One rank of the actor-learner loop.

    Two copies of the network, like the reference, where the workers play with the weights of the last finished
    epoch while the trainer moves on (worker.py:41-56, train.py:403-414): `actor` (eval mode) drives generation,
    `net` is trained.  With `overlap` the generation wave of epoch e runs on its own CUDA stream BESIDE the training
    steps + gradient all-reduces of epoch e (which read windows of the episodes finished up to wave e-1) - SURVEY
    8 E2: collectives on a side stream overlapped with the next rollout wave.  At the epoch boundary the learner's
    weights are broadcast from rank 0 (one flat NCCL broadcast) and copied into the actor.

    What keeps the overlap race-free: the trainer's window selection reads a SNAPSHOT of the episode-table counters
    taken on the generation stream at the end of the previous wave (rows below that cursor are complete), and
    skips episodes whose first ply the running wave could overwrite in the ring (`it_now` = the iteration counter
    at the END of the running wave)."""

    def __init__(self, env_name, n_games, args, device, rank=0, world=1, overlap=True):
        self.args, self.device, self.rank, self.world = args, device, rank, world
        if env_name == 'ParallelTicTacToe':
            args['turn_based_training'] = False          # simultaneous game: one player per episode (train.py:57-58)
        torch.manual_seed(args['seed'])
        self.net = make_net(env_name).to(device)
        self.actor = make_net(env_name).to(device)
        self.actor.eval()
        depth = 1024 if 'Geister' in env_name else 256
        self.gen = BatchGenerator(env_name, n_games, device=device, seed=args['seed'], game_id_base=rank * n_games,
                                  depth=depth, max_episodes=1 << 20, record_values=True, record_logits=True)
        self.learner = DataParallelLearner(self.net, args)          # broadcasts rank 0's initial weights
        self._sync_actor()
        self.n_games = n_games
        self.hidden = self.gen.init_hidden(self.actor)      # per player: (n_games, 2, ...)
        self.batch = None
        self.overlap = overlap
        self.gen_stream = torch.cuda.Stream(device=device)
        self.train_stream = torch.cuda.Stream(device=device)
        self.snap = torch.zeros(8, dtype=torch.int64, device=device)   # counters at the end of the last finished wave
        self.windows = torch.empty(args['batch_size'], 8, dtype=torch.int32, device=device)
        self.ev_wave = torch.cuda.Event()
        self.ev_train = torch.cuda.Event()
        self.timing = {'allreduce_ms': [], 'broadcast_ms': []}

    def _init_hidden(self, lead):
        if not hasattr(self.net, 'init_hidden'):
            return None
        h = self.net.init_hidden(lead)
        return tuple([t.to(self.device) for t in part] for part in h)

    @torch.no_grad()
    def _sync_actor(self):
        for dst, src in zip(list(self.actor.parameters()) + list(self.actor.buffers()),
                            list(self.net.parameters()) + list(self.net.buffers())):
            dst.copy_(src)

    @torch.no_grad()
    def generate(self, plies):
        """one wave on the generation stream; afterwards the counter snapshot of the finished wave"""
        with torch.cuda.stream(self.gen_stream):
            self.hidden = self.gen.play(self.actor, plies, self.hidden)
            self.snap_next = self.gen.store.counters.clone()
            self.ev_wave.record()

    def train(self, steps, it_now):
        a = self.args
        stats = None
        with torch.cuda.stream(self.train_stream):
            for _ in range(steps):
                w = self.gen.select_windows(a['batch_size'], a, out=self.windows, counters=self.snap, it_now=it_now)
                self.batch = make_batch_windows(self.gen.store, w, a, out=self.batch)
                B, P = self.batch['value'].shape[0], self.batch['value'].shape[2]
                stats = self.learner.train_step(self.batch, self._init_hidden([B, P]))
            self.learner.end_epoch()
            self.ev_train.record()
        return stats

    def _ready(self):
        """train.py:330-331: no training before `minimum_episodes` - decided COLLECTIVELY (every rank must enter the
        same sequence of NCCL calls): the minimum over the ranks of the episodes usable by the trainer."""
        return all_ranks_have(self.snap[0:1], self.args['minimum_episodes'])

    def epoch(self, plies, train_steps):
        """generation wave e  ||  training on the episodes of waves < e;  then weights: rank 0 -> all ranks -> actor"""
        cur = torch.cuda.current_stream()
        self.gen_stream.wait_stream(cur)
        self.train_stream.wait_stream(cur)
        ready = self._ready()                                # uses the snapshot of the previous wave
        self.generate(plies)                                 # enqueued on gen_stream, returns at once
        if not self.overlap:
            self.train_stream.wait_event(self.ev_wave)
        out = None
        if ready:
            out = self.train(train_steps, it_now=self.gen.store.it)
        with torch.cuda.stream(self.train_stream):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            broadcast_weights(self.net, 0)                   # train.py:613 / worker.py:52 stand-in: one flat broadcast
            e1.record()
            self.train_stream.wait_event(self.ev_wave)       # the actor is in use until the wave is done
            self._sync_actor()
            self.snap.copy_(self.snap_next)
        cur.wait_stream(self.train_stream)
        cur.wait_stream(self.gen_stream)
        self._bcast_events = (e0, e1)
        return out

    def collective_times(self):
        """CUDA-event milliseconds of the last flat broadcast and of one flat gradient all-reduce (measured alone)"""
        torch.cuda.synchronize()
        out = {'broadcast_ms': self._bcast_events[0].elapsed_time(self._bcast_events[1])}
        if self.world > 1:
            from .learner import flat_grads
            params = [p for p in self.net.parameters() if p.requires_grad]
            bucket = torch.zeros(sum(p.numel() for p in params) + 1, device=self.device)
            for _ in range(3):
                dist.all_reduce(bucket)
            ts = []
            for _ in range(10):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dist.all_reduce(bucket)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            out['allreduce_ms'] = ts[len(ts) // 2]
            out['allreduce_bytes'] = bucket.numel() * 4
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
    ap.add_argument('--minimum-episodes', type=int, default=400)
    ap.add_argument('--no-overlap', action='store_true')
    a = ap.parse_args()
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    args = dict(DEFAULT_ARGS, batch_size=a.batch_size, value_target=a.value_target, policy_target=a.policy_target,
                minimum_episodes=a.minimum_episodes)
    al = ActorLearner(a.env, a.games, args, dev, rank, world, overlap=not a.no_overlap)
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
