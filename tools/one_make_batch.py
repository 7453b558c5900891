"""A few make_batch launches of one game (for ncu):  python tools/one_make_batch.py GAME [T] [B]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from handyrl_b200.generation import BatchGenerator
from handyrl_b200.train import make_batch_windows

game = sys.argv[1]
T = int(sys.argv[2]) if len(sys.argv) > 2 else (32 if 'Geister' in game else 8)
N = 65536
B = int(sys.argv[3]) if len(sys.argv) > 3 else N * 32 // T
args = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': T, 'compress_steps': 4,
        'maximum_episodes': 1 << 20, 'lambda': 0.7}
gen = BatchGenerator(game, N, seed=1, depth=512, max_episodes=1 << 21)
for _ in range(8):
    gen.rollout(64)
win = gen.select_windows(B, args)
out = make_batch_windows(gen.store, win, args)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(4):
    if i == 3:
        e0.record()
    make_batch_windows(gen.store, win, args, out=out)
e1.record()
torch.cuda.synchronize()
print(game, B, T, "%.4f ms" % e0.elapsed_time(e1))
