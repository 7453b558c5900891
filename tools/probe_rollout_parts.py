"""What each side output of k_rollout costs on the bench workload (65,536 Geister games x 32 plies per launch):
full = moments + action + trajectory records + episode table; then without the record ring, without the episode
table, with the static schedule (knob rollout_chunk = 0), with bulk-store emit (knob emit_tma = 1).
    python tools/probe_rollout_parts.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200 import _lib  # noqa: E402
from handyrl_b200.engine import BatchedEnv  # noqa: E402

N, PLIES, D = 65536, 32, 512


def run(name, traj=True, table=True, knobs=(), D=D):
    for k, v in knobs:
        _lib.set_tuning(k, v)
    env = BatchedEnv("Geister", N)
    main, scalar = env.alloc_obs(waves=2)
    waves = dict(main=main, scalar=scalar, action_mask=torch.empty(2, N, env.A, device="cuda"))
    act = torch.empty(2, N, dtype=torch.int32, device="cuda")
    tr = torch.zeros(N, D, 8, dtype=torch.int32, device="cuda") if traj else None
    ept = torch.zeros(1 << 20, 4, dtype=torch.int32, device="cuda") if table else None
    ctr = torch.zeros(8, dtype=torch.int64, device="cuda")
    for _ in range(10):
        env.rollout(PLIES, seed=1, waves=waves, traj=tr, ep_table=ept, counters=ctr, out_action=act)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K = 200
    for _ in range(K):
        env.rollout(PLIES, seed=1, waves=waves, traj=tr, ep_table=ept, counters=ctr, out_action=act)
    e1.record()
    torch.cuda.synchronize()
    for k, _ in knobs:
        _lib.set_tuning(k, None)
    ms = e0.elapsed_time(e1) / K
    byt = N * PLIES * (1940 + (32 if traj else 0))
    return {"case": name, "ms": round(ms, 4), "TBps": round(byt / ms / 1e9, 3)}


if __name__ == "__main__":
    rows = [run("full"), run("no trajectory ring", traj=False), run("no episode table", table=False),
            run("no ring, no table", traj=False, table=False),
            run("static schedule", knobs=(("rollout_chunk", 0),)), run("bulk-store emit", knobs=(("emit_tma", 1),)),
            run("tickets of 8 plies", knobs=(("rollout_chunk", 8),)),
            run("tickets of 16 plies", knobs=(("rollout_chunk", 16),)),
            run("tickets of 8 plies, no ring", traj=False, knobs=(("rollout_chunk", 8),)),
            # ring depth: 512 makes the slot stride a power of two (16 KB): every game's record line of a ply group
            # has the same low 14 address bits
            run("ring depth 516", D=516), run("ring depth 520", D=520), run("ring depth 544", D=544),
            run("ring depth 772", D=772),
            # tiny rings: the same instructions, but the 32 lines a warp flushes are (nearly) contiguous
            run("ring depth 4 (4 KB contiguous per warp flush)", D=4), run("ring depth 8", D=8),
            run("ring depth 32", D=32), run("ring depth 128", D=128),
            run("full (again)")]
    for r in rows:
        print(r, file=sys.stderr)
    print(json.dumps(rows))
