"""One target-recursion shape, a few launches (for ncu):  python tools/one_target.py MODE B T [alias]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from handyrl_b200.losses import compute_target, compute_target_pair

mode, B, T = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
alias = len(sys.argv) > 4
P = 2
v = torch.tanh(torch.randn(B, T, P, 1, device='cuda'))
ret = torch.randn(B, T, P, 1, device='cuda')
rew = torch.full((B, T, P, 1), -0.01, device='cuda')
rho = torch.rand(B, T, 1, 1, device='cuda')
c = rho if alias else rho.clone()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(4):
    if i == 3:
        e0.record()
    if mode == 'PAIR':
        compute_target_pair('VTRACE', 'UPGO', v, ret, rew, 0.7, 0.8, rho, c)
    else:
        compute_target(mode, v, ret, rew, 0.7, 0.8, rho, c)
e1.record()
torch.cuda.synchronize()
print(mode, B, T, "%.4f ms" % e0.elapsed_time(e1))
