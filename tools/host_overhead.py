"""Host-side cost per call of the Python shims (wall clock over many tiny launches)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200.losses import compute_target, compute_target_pair  # noqa: E402

dev = torch.device("cuda")
B, T, P = 64, 16, 2
v = torch.randn(B, T, P, 1, device=dev)
ret = torch.randn(B, T, P, 1, device=dev)
rew = torch.randn(B, T, P, 1, device=dev)
oc = torch.randn(B, 1, P, 1, device=dev)
rho = torch.rand(B, T, 1, 1, device=dev)
c = rho.clone()
for name, fn in [("compute_target VTRACE", lambda: compute_target("VTRACE", v, ret, rew, 0.7, 0.8, rho, c)),
                 ("compute_target_pair UPGO/VTRACE value head", lambda: compute_target_pair("UPGO", "VTRACE", v, oc, None, 0.7, 1.0, rho, c)),
                 ("compute_target MC", lambda: compute_target("MC", v, oc, None, 0.7, 1.0, rho, c)),
                 ("torch.empty_like x2 (reference point)", lambda: (torch.empty_like(v), torch.empty_like(v)))]:
    for _ in range(200):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 3000
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    print("%-50s %.1f us/call" % (name, (time.perf_counter() - t0) / n * 1e6))
