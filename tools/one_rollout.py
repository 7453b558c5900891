"""A few rollout launches of one game (for ncu):  python tools/one_rollout.py GAME [N] [PLIES]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from handyrl_b200.generation import BatchGenerator

game = sys.argv[1]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
plies = int(sys.argv[3]) if len(sys.argv) > 3 else (32 if 'Geister' in game else 256)
gen = BatchGenerator(game, N, seed=1, depth=512, max_episodes=1 << 20)
env = gen.env
W = 2 if 'Geister' in game else 32
main_w, scalar_w = env.alloc_obs(waves=W)
waves = dict(main=main_w, scalar=scalar_w, action_mask=torch.empty(W, N, env.A, device='cuda'))
act = torch.empty(W, N, dtype=torch.int32, device='cuda')
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(4):
    if i == 3:
        e0.record()
    gen.rollout(plies, waves=waves, out_action=act)
e1.record()
torch.cuda.synchronize()
print(game, N, plies, "%.4f ms" % e0.elapsed_time(e1))
