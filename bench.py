#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native HandyRL hot path.

Workload (BASELINE.json configs[1]): Geister self-play, 65,536 concurrent games per GPU,
epoch-0 policy (uniform over legal actions == RandomModel), V-Trace policy + UPGO value
targets.  One *step* is one pass of the hot path over one batch of synthetic input:

  1. rollout   : PLIES plies for all 65,536 games in one persistent launch; every ply
                 writes the reference-format moment (observation 7x6x6 + 18, action_mask
                 214 fp32, action) plus a 32-byte trajectory record          -> env-steps
  2. make_batch: B windows of T plies gathered from the device trajectory ring into the
                 (B,T,P,...) training batch (when the kernel is available)
  3. targets   : compute_target pair (UPGO value targets + V-Trace policy advantages) for
                 the value head and the return head on (B,T,2,1)              -> target timesteps

`value` = env-steps/s of the whole job (all ranks) over the wall time of the K timed steps
(device time, max over ranks).  Inputs are device resident; `e2e` repeats the stepping
through host buffers (pinned H2D actions, D2H observation/mask/reward/terminal).

  python bench.py --gpus N --steps K --warmup W          # N>1 under torchrun
  python bench.py --impl reference ...                   # CPU oracle port on the host cores
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_GAMES = 65536
PLIES = 32
B, T, P = 65536, 32, 2
BYTES_PER_ENV_STEP = 1940          # obs 1008 + scalar 72 + action_mask 856 + action 4
BYTES_PER_TARGET_ELEM = 20         # fused V-Trace + UPGO pass: v, r, rho/2, c/2 in; target, adv out


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, reasons, mx = [], set(), None
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx = float(f[2])
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                   f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))
        return out


def cpu_baseline(seconds_target=12.0):
    """The C oracle's self-play loop (oracle/games_oracle.c: orc_selfplay) on all host cores."""
    from oracle import games_c
    cores = os.cpu_count() or 1
    # calibrate, then run a bounded sample
    t0 = time.perf_counter()
    n = games_c.selfplay("geister", 20000, seed=1, n_threads=cores)
    dt = time.perf_counter() - t0
    rate = n / dt
    per_thread = max(20000, int(rate * seconds_target / cores))
    t0 = time.perf_counter()
    n = games_c.selfplay("geister", per_thread, seed=2, n_threads=cores)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": "%d Geister env-steps (obs + action_mask + policy + uniform action + play) per thread x %d "
                      "pthreads of the C oracle, %.1f s" % (per_thread, cores, dt)}


def run_reference(args, rank, world):
    """--impl reference: the reference is pure Python and cannot travel to the GPU box; the timed arm is
    the CPU oracle port (C restatement of the same per-ply work) on all host cores."""
    if rank != 0:
        return
    from oracle import games_c
    cores = os.cpu_count() or 1
    per_thread = 150000
    for _ in range(args.warmup):
        games_c.selfplay("geister", 20000, seed=0, n_threads=cores)
    t0 = time.perf_counter()
    total = 0
    for k in range(args.steps):
        total += games_c.selfplay("geister", per_thread, seed=10 + k, n_threads=cores)
    dt = time.perf_counter() - t0
    v = total / dt
    line = {
        "impl": "reference", "metric": "geister_selfplay_env_steps_per_sec", "value": v, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 bitboards -> f32",
        "data": "synthetic", "config": workload_config(args.gpus),
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                         "sample": "%d env-steps per thread x %d pthreads per step (C oracle orc_selfplay)"
                                   % (per_thread, cores)},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus):
    return {"workload": "Geister self-play, %d concurrent games/GPU x %d plies per step (uniform epoch-0 policy, "
                        "auto-reset), full moment output per ply; then V-Trace policy + UPGO value targets on "
                        "(B=%d,T=%d,P=%d) for value and return heads" % (N_GAMES, PLIES, B, T, P),
            "games_per_gpu": N_GAMES, "plies_per_step": PLIES, "target_shape": [B, T, P],
            "value_target": "UPGO", "policy_target": "VTRACE", "lambda": 0.7, "gamma": 0.8,
            "l2_policy": "inputs larger than L2: 2 waves x 127 MB of moment output, 4 x 16.8 MB target inputs "
                         "rotated over 8 buffer sets",
            "parallelism": "games sharded per GPU, no data-path collective (dp%d)" % n_gpus}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from handyrl_b200.engine import BatchedEnv
    from handyrl_b200.losses import compute_target_pair

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---------------- device-resident workload ----------------
    env = BatchedEnv("Geister", N_GAMES, device=dev)
    WAVES = 2
    main_w, scalar_w = env.alloc_obs(waves=WAVES)
    waves = dict(main=main_w, scalar=scalar_w, action_mask=torch.empty(WAVES, N_GAMES, env.A, device=dev))
    act_w = torch.empty(WAVES, N_GAMES, dtype=torch.int32, device=dev)
    D = 512
    traj = torch.zeros(D, 2, N_GAMES, 4, dtype=torch.int32, device=dev)
    ept = torch.zeros(1 << 20, 4, dtype=torch.int32, device=dev)
    ctr = torch.zeros(8, dtype=torch.int64, device=dev)

    g = torch.Generator(device=dev).manual_seed(rank)
    NSETS = 8  # rotate input sets so that consecutive steps do not hit L2 (8 x 4 x 16.8 MB > 126 MB)
    tsets = []
    for _ in range(NSETS):
        v = torch.tanh(torch.randn(B, T, P, 1, device=dev, generator=g))
        rho = torch.exp(0.5 * torch.randn(B, T, 1, 1, device=dev, generator=g)).clamp(0, 1)
        oc = torch.randint(-1, 2, (B, 1, P, 1), device=dev, generator=g).float()
        rew = torch.full((B, T, P, 1), -0.01, device=dev)
        ret = torch.randn(B, T, P, 1, device=dev, generator=g)
        vr = torch.tanh(torch.randn(B, T, P, 1, device=dev, generator=g))
        tsets.append((v, oc, rho, vr, ret, rew))

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def step(k, evs=None):
        if evs:
            evs[0].record()
        env.rollout(PLIES, seed=1234, game_id_base=rank * N_GAMES, waves=waves, traj=traj, ep_table=ept,
                    counters=ctr, out_action=act_w)
        if evs:
            evs[1].record()
        v, oc, rho, vr, ret, rew = tsets[k % NSETS]
        compute_target_pair("UPGO", "VTRACE", v, oc, None, 0.7, 1.0, rho, rho)
        compute_target_pair("UPGO", "VTRACE", vr, ret, rew, 0.7, 0.8, rho, rho)
        if evs:
            evs[2].record()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for k in range(args.warmup):
        step(k)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    phase_evs = [[ev(), ev(), ev()] for _ in range(args.steps)]
    e0, e1 = ev(), ev()
    barrier()
    e0.record()
    for k in range(args.steps):
        step(k, phase_evs[k])
    e1.record()
    barrier()
    total_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    roll_ms = sum(p[0].elapsed_time(p[1]) for p in phase_evs) / args.steps
    tgt_ms = sum(p[1].elapsed_time(p[2]) for p in phase_evs) / args.steps
    if world > 1:
        tt = torch.tensor([total_ms, roll_ms, tgt_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, roll_ms, tgt_ms = [float(x) for x in tt.tolist()]

    env_steps_per_step = N_GAMES * PLIES * world
    value = env_steps_per_step * args.steps / (total_ms * 1e-3)
    tgt_elems = 2 * B * T * P * world

    # ---------------- e2e: host buffers through the public API ----------------
    e2e = None
    if not args.no_e2e:
        e2e_env = BatchedEnv("Geister", N_GAMES, device=dev)
        out = e2e_env.step_observe(None)
        h_act = torch.zeros(N_GAMES, dtype=torch.int32).pin_memory()
        h = {k: torch.empty(t.shape, dtype=t.dtype).pin_memory() for k, t in out.items() if t is not None}
        d_act = torch.empty(N_GAMES, dtype=torch.int32, device=dev)
        E2E_PLIES = 8

        def e2e_step():
            nonlocal out
            for _ in range(E2E_PLIES):
                # host policy on the host copy of the mask: first legal action (argmin of the 0 / 1e32 mask)
                torch.argmin(h["action_mask"], dim=1, out=h_idx)
                h_act.copy_(h_idx)
                d_act.copy_(h_act, non_blocking=True)                      # H2D: this ply's actions
                out = e2e_env.step_observe(d_act, auto_reset=True, out=out)
                for k2, t in h.items():                                     # D2H: obs, mask, reward, terminal ...
                    t.copy_(out[k2], non_blocking=True)
                torch.cuda.current_stream().synchronize()

        h_idx = torch.zeros(N_GAMES, dtype=torch.int64)
        for k2, t in h.items():
            t.copy_(out[k2])
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(world, 1)))
        for _ in range(2):
            e2e_step()
        barrier()
        ksteps = max(2, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(ksteps):
            e2e_step()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        d2h = sum(t.numel() * t.element_size() for t in h.values())
        e2e = {"value": N_GAMES * E2E_PLIES * ksteps * world / dt, "unit": "env-steps/s",
               "h2d_bytes_per_step": int(h_act.numel() * 4 * E2E_PLIES), "d2h_bytes_per_step": int(d2h * E2E_PLIES),
               "plies_per_step": E2E_PLIES,
               "note": "BatchedEnv.step_observe per ply with pinned-host actions in and observation/action_mask/"
                       "reward/terminal/outcome/turn out; PCIe-bound"}

    if rank == 0:
        peak, peak_kind = load_peaks()
        achieved = BYTES_PER_ENV_STEP * N_GAMES * PLIES / (roll_ms * 1e-3) / 1e9
        t_ach = BYTES_PER_TARGET_ELEM * (tgt_elems / world) / (tgt_ms * 1e-3) / 1e9
        line = {
            "metric": "geister_selfplay_env_steps_per_sec", "value": value, "unit": "env-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64 bitboards -> f32", "data": "synthetic", "config": workload_config(world),
            "target_timesteps_per_sec": tgt_elems / (tgt_ms * 1e-3),
            "phases": {"rollout_ms": roll_ms, "targets_ms": tgt_ms,
                       "env_steps_per_sec_rollout_only": env_steps_per_step / (roll_ms * 1e-3)},
            "roofline": {"kernel": "k_rollout<Geister>", "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_kind": peak_kind,
                         "bytes_per_unit": BYTES_PER_ENV_STEP},
            "roofline_targets": {"kernel": "k_target_scan<UPGO,VTRACE>", "bound": "hbm", "achieved": t_ach,
                                 "peak": peak, "unit": "GB/s", "frac": t_ach / peak,
                                 "bytes_per_unit": BYTES_PER_TARGET_ELEM},
            "clocks": clocks, "e2e": e2e, "gpu_launches": 3 * args.steps,
        }
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
