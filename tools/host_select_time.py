"""Wall time of HostReplayIndex.select (b200rl_select_windows_host) by thread count, after a few rollouts."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from handyrl_b200.generation import BatchGenerator, HostReplayIndex

ARGS = {'maximum_episodes': 1 << 20, 'forward_steps': 32}
gen = BatchGenerator("Geister", 65536, depth=1024, max_episodes=1 << 20)
for _ in range(12):
    gen.rollout(32)
torch.cuda.synchronize()
print("cpu_count", os.cpu_count())
for nt in (1, 2, 4, 8, 16, 32):
    idx = HostReplayIndex(gen, n_threads=nt)
    buf = torch.empty(65536, 8, dtype=torch.int32).pin_memory()
    idx.select(65536, ARGS, buf)
    t0 = time.perf_counter()
    for _ in range(10):
        idx.select(65536, ARGS, buf)
    print("threads %2d: %.3f ms per 65536 windows" % (nt, (time.perf_counter() - t0) / 10 * 1e3))
idx = HostReplayIndex(gen, n_threads=8)
t0 = time.perf_counter()
for _ in range(10):
    gen.rollout(32)
    idx.fetch_async()
    idx.commit()
print("rollout + fetch + commit: %.3f ms" % ((time.perf_counter() - t0) / 10 * 1e3))
