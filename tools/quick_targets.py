"""Quick device timing of the target scan (CUDA events), for development."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from handyrl_b200.losses import compute_target, compute_target_pair

BYTES = {'MC': 12, 'TD': 16, 'UPGO': 16, 'VTRACE': 20}
for (B, T) in [(65536, 32), (4096, 32), (4096, 200), (65536, 200), (4096, 1024), (65536, 1024)]:
    P = 2
    v = torch.tanh(torch.randn(B, T, P, 1, device='cuda'))
    ret = torch.randn(B, T, P, 1, device='cuda')
    rew = torch.full((B, T, P, 1), -0.01, device='cuda')
    rho = torch.rand(B, T, 1, 1, device='cuda')
    for mode in ['MC', 'TD', 'UPGO', 'VTRACE', 'VT+UPGO']:
        def run():
            if mode == 'VT+UPGO':
                return compute_target_pair('VTRACE', 'UPGO', v, ret, rew, 0.7, 0.8, rho, rho)
            return compute_target(mode, v, ret, rew, 0.7, 0.8, rho, rho)
        for _ in range(3):
            run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        el = B * T * P
        by = BYTES.get(mode, 20)
        print("B=%6d T=%4d %-8s %8.3f ms  %8.2f Gelem/s  %7.1f GB/s" % (B, T, mode, ms, el / ms / 1e6, el * by / ms / 1e6), flush=True)
