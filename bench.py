#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native HandyRL hot path.

Workload (BASELINE.json configs[1]): Geister self-play, 65,536 concurrent games per GPU,
epoch-0 policy (uniform over legal actions == RandomModel), V-Trace policy + UPGO value
targets.  One *step* is one pass of the whole hot path over one batch of synthetic input:

  1. rollout     : PLIES plies for all 65,536 games in one persistent launch; every ply writes the
                   reference-format moment (observation 7x6x6 + 18, action_mask 214 fp32, action)
                   plus a 32-byte trajectory record                              -> env-steps
  2. select      : B training windows drawn on the device from the finished-episode table
                   (Batcher.select_episode, train.py:284-302)
  3. make_batch  : the B windows x T plies re-encoded from the trajectory ring into the
                   reference's 12-tensor (B,T,P,...) batch (train.py:33-133)      -> batch timesteps
  4. targets     : compute_target pair (UPGO value targets + V-Trace policy advantages) for the
                   value head (bootstrap = batch outcome) and the return head (batch return /
                   reward) on (B,T,2,1)                                          -> target timesteps

`value` = env-steps/s of the whole job (all ranks) over the device time of the K timed steps (max
over ranks); every generated ply is batched and targeted once (B*T == games*plies).  Inputs are
device resident.  `e2e` runs the same step through the public API with the replay index on the
HOST: finished-episode rows are read back, windows are selected with numpy and uploaded from pinned
memory, and a scalar summary of the targets is read back every step.

  python bench.py --gpus N --steps K --warmup W          # N>1 under torchrun
  python bench.py --impl reference ...                   # CPU oracle port on the host cores
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_GAMES = 65536
PLIES = 32
B, T, P = 65536, 32, 2
RING_DEPTH = 512
TRAIN_ARGS = {"turn_based_training": True, "observation": False, "gamma": 0.8, "forward_steps": T,
              "compress_steps": 4, "maximum_episodes": 100000, "lambda": 0.7,
              "policy_target": "VTRACE", "value_target": "UPGO"}
# Algorithmic HBM bytes per unit
BYTES_PER_ENV_STEP = 1940 + 32     # obs 1008 + scalar 72 + action_mask 856 + action 4, + 32 B trajectory record
BYTES_PER_BATCH_CELL = 2848 + 32   # the 12 batch tensors of one (b,t) in TURN mode, + the 32 B record read
BYTES_PER_TARGET_PAIR = 16 + 20    # value head (v, rho/2, c/2 in; 2 out) + return head (+ rewards), per (b,t,p)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


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
                 "-lms", "20"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
            time.sleep(0.3)  # let the first samples land before the timed region starts
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.1)
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


def workload_config(n_gpus):
    return {"workload": "Geister self-play hot path per step: %d concurrent games/GPU x %d plies (uniform epoch-0 "
                        "policy, auto-reset, full fp32 moment per ply) -> device select_episode -> make_batch "
                        "(B=%d,T=%d, turn-based) -> UPGO value targets + V-Trace policy advantages for the value and "
                        "return heads on (B,T,%d,1)" % (N_GAMES, PLIES, B, T, P),
            "games_per_gpu": N_GAMES, "plies_per_step": PLIES, "batch_shape": [B, T, P],
            "value_target": "UPGO", "policy_target": "VTRACE", "lambda": 0.7, "gamma": 0.8,
            "l2_policy": "inputs larger than L2: each step streams 4.1 GB of moments, 6.0 GB of batch tensors and "
                         "rotates 8 sets of synthetic net outputs (4 x 16.8 MB each) against the 126 MB L2",
            "parallelism": "games, windows and target rows sharded per GPU, no data-path collective (dp%d)" % n_gpus}


# ---------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the C oracle port on the host cores
# ---------------------------------------------------------------------------------------------
CPU_GAMES_PER_THREAD = 256


def cpu_hotpath_runner(seed):
    """The whole hot path on all host cores through the C oracle port (oracle/hotpath_oracle.c): per thread
    256 Geister games x PLIES plies -> select_episode + make_batch (B = 256*PLIES/T windows) -> UPGO/V-Trace
    targets for both heads.  Returns (runner, cores) with the replay lists pre-filled (untimed)."""
    from oracle import games_c
    cores = os.cpu_count() or 1
    hp = games_c.HotPath(CPU_GAMES_PER_THREAD, PLIES, T, cores, replay_cap=1024, seed=seed,
                         gamma=TRAIN_ARGS['gamma'], lmb=TRAIN_ARGS['lambda'])
    for _ in range(8):      # Geister games last ~157 plies: let every thread finish episodes before timing
        hp.step()
    return hp, cores


def cpu_sample_text(cores, steps, dt):
    return ("%d steps x %d pthreads x (%d Geister games x %d plies -> select_episode + make_batch of %d windows x "
            "T=%d -> UPGO targets + V-Trace advantages, value and return heads), C oracle port, no bz2/pickle, %.1f s"
            % (steps, cores, CPU_GAMES_PER_THREAD, PLIES, CPU_GAMES_PER_THREAD * PLIES // T, T, dt))


def cpu_baseline(seconds_target=12.0):
    hp, cores = cpu_hotpath_runner(1)
    t0 = time.perf_counter()
    n = hp.step()[0]
    per_step = time.perf_counter() - t0
    steps = max(3, int(seconds_target / max(per_step, 1e-3)))
    p0 = hp.phase_seconds()
    t0 = time.perf_counter()
    n = sum(hp.step()[0] for _ in range(steps))
    dt = time.perf_counter() - t0
    p1 = hp.phase_seconds()
    ph = [(y - x) for x, y in zip(p0, p1)]
    tot = sum(ph) or 1.0
    return {"value": n / dt, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": cpu_sample_text(cores, steps, dt),
            "phase_share": {"generation": ph[0] / tot, "make_batch": ph[1] / tot, "targets": ph[2] / tot}}


def cpu_baseline_reference(seconds=6.0, parts="generation,generation_net,make_batch,targets"):
    """The UNMODIFIED reference's own Python path (baseline/_ref, installed by tools/install_ref.sh) timed on this
    box's host cores in os.cpu_count() single-threaded processes: Generator.generate (RandomModel and GeisterNet),
    make_batch, losses.compute_target - a separate process so that its torch thread settings and
    forks stay away from this one."""
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "time_reference.py"), "--seconds", str(seconds),
                            "--parts", parts], stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, timeout=300, text=True)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as ex:       # reported, never fatal for the GPU numbers
        return {"unavailable": "time_reference.py failed: %r" % (ex,)}


def learner_leg(dev, rank, world, barrier):
    """BASELINE configs[2] and [3]: the actor-learner loop (handyrl_b200/loop.py) with the network in the loop - the
    only place where this path has collectives: one flat gradient all-reduce per optimiser step and one flat weight
    broadcast per epoch (NCCL), on the training stream beside the next generation wave.  Reported next to the
    device-only numbers, never mixed into `value`: >99 % of a ply here is the PyTorch forward (library code)."""
    import torch
    import torch.distributed as dist
    from handyrl_b200.loop import ActorLearner, DEFAULT_ARGS
    out = {}
    cases = (("Geister", 16384, 16, "VTRACE", "UPGO", 10, 192), ("ParallelTicTacToe", 65536, 16, "TD", "TD", 10, 32))
    for env_name, games, plies, vt, pt, epochs, prefill in cases:
        args = dict(DEFAULT_ARGS, value_target=vt, policy_target=pt, batch_size=256, minimum_episodes=400)
        al = ActorLearner(env_name, games, args, dev, rank, world, overlap=True)
        al.gen.rollout(prefill)                      # epoch-0 policy (zero logits) fills the replay before timing
        al.snap.copy_(al.gen.store.counters)
        train_steps = 4
        for _ in range(2):
            res = al.epoch(plies, train_steps)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(epochs):
            res = al.epoch(plies, train_steps)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            tt = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        coll = al.collective_times()
        out[env_name] = {
            "games_per_gpu": games, "plies_per_epoch": plies, "train_steps_per_epoch": train_steps,
            "batch": [args["batch_size"], args["forward_steps"]], "value_target": vt, "policy_target": pt,
            "epochs_timed": epochs, "ms_per_epoch": ms / epochs,
            "env_steps_per_s": games * plies * world * epochs / (ms * 1e-3),
            "train_steps_per_s": train_steps * epochs / (ms * 1e-3),
            "trained": res is not None, "loss_total": (res[0]["total"] if res else None),
            "net": type(al.net).__name__, "params": sum(p.numel() for p in al.net.parameters()),
            "collectives": coll, "overlap": "generation stream || training stream (+NCCL)", "ranks": world}
        del al
        torch.cuda.empty_cache()
    return out


RESULT_OUT = sys.stdout


def run_reference(args, rank, world):
    """--impl reference: the reference is pure Python and cannot travel to the GPU box (in the build container it
    runs ~5.4 k Geister plies/s/process); the timed arm is the C oracle port of the same step on
    all host cores."""
    if rank != 0:
        return
    hp, cores = cpu_hotpath_runner(0)
    # one reference step = the N=1 workload of one step (N_GAMES games x PLIES plies), as `inner` rounds of
    # cores x 256 games: ~0.3 s per step on 16 cores, so that K = 20 steps are a 6 s sample and not 0.4 s
    inner = max(1, N_GAMES // (cores * CPU_GAMES_PER_THREAD))
    for _ in range(args.warmup):
        for _ in range(inner):
            hp.step()
    t0 = time.perf_counter()
    total = 0
    for k in range(args.steps):
        for _ in range(inner):
            total += hp.step()[0]
    dt = time.perf_counter() - t0
    v = total / dt
    line = {
        "impl": "reference", "metric": "geister_selfplay_env_steps_per_sec", "value": v, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32 cell arrays -> f32",
        "data": "synthetic", "config": workload_config(args.gpus),
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                         "sample": cpu_sample_text(cores, args.steps * inner, dt)},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=RESULT_OUT, flush=True)


# ---------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-learner", action="store_true")
    ap.add_argument("--rounds", type=int, default=0, help="repeat the K timed steps this many times (0: >= 0.35 s)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    # stdout carries exactly ONE line (the JSON): library chatter written to fd 1 (NCCL prints its version banner
    # there) goes to stderr, the result line to a private duplicate of the original stdout
    global RESULT_OUT
    sys.stdout.flush()
    RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from handyrl_b200.generation import BatchGenerator, HostReplayIndex
    from handyrl_b200.losses import compute_target_heads
    from handyrl_b200.train import make_batch_windows

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---------------- device-resident workload ----------------
    gen = BatchGenerator("Geister", N_GAMES, device=dev, seed=1234, game_id_base=rank * N_GAMES, depth=RING_DEPTH,
                         max_episodes=1 << 20, args={})
    env = gen.env
    WAVES = 2
    main_w, scalar_w = env.alloc_obs(waves=WAVES)
    waves = dict(main=main_w, scalar=scalar_w, action_mask=torch.empty(WAVES, N_GAMES, env.A, device=dev))
    act_w = torch.empty(WAVES, N_GAMES, dtype=torch.int32, device=dev)
    windows = torch.empty(B, 8, dtype=torch.int32, device=dev)

    g = torch.Generator(device=dev).manual_seed(rank)
    NSETS = 8  # rotate the synthetic net outputs so that consecutive steps do not hit L2
    tsets = []
    for _ in range(NSETS):
        v = torch.tanh(torch.randn(B, T, P, 1, device=dev, generator=g))        # value head output
        vr = torch.tanh(torch.randn(B, T, P, 1, device=dev, generator=g))       # return head output
        rho = torch.exp(0.5 * torch.randn(B, T, 1, 1, device=dev, generator=g)).clamp(0, 1)   # train.py:230
        c = rho.clone()                                                          # train.py:231
        tsets.append((v, vr, rho, c))

    ev = lambda: torch.cuda.Event(enable_timing=True)
    state = {"batch": None}

    def step(k, evs=None, win=None):
        rec = (lambda i: evs[i].record()) if evs else (lambda i: None)
        rec(0)
        gen.rollout(PLIES, waves=waves, out_action=act_w)
        rec(1)
        if win is None:
            gen.select_windows(B, TRAIN_ARGS, out=windows)
            win = windows
        state["batch"] = batch = make_batch_windows(gen.store, win, TRAIN_ARGS, out=state["batch"])
        rec(2)
        v, vr, rho, c = tsets[k % NSETS]
        # train.py:245-253 with value_target UPGO / policy_target VTRACE: targets of the first, advantages of the second
        t1, t2 = compute_target_heads("UPGO", "VTRACE", [(v, batch["outcome"], None, 1.0),
                                                          (vr, batch["return"], batch["reward"], 0.8)], 0.7, rho, c)
        rec(3)
        return t1, t2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # fill the finished-episode table before anything is timed (Geister games last ~157 plies)
    for _ in range(8):
        gen.rollout(PLIES, waves=waves, out_action=act_w)
    for k in range(args.warmup):
        step(k)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # The K timed steps are repeated in `rounds` back-to-back rounds so that the timed region lasts >= 0.35 s (20
    # steps are only 33 ms: one or two nvidia-smi samples); every number below is the mean over K * rounds steps.
    t_probe0, t_probe1 = ev(), ev()
    t_probe0.record()
    step(0)
    t_probe1.record()
    torch.cuda.synchronize()
    rounds = args.rounds or max(1, int(0.35 / (args.steps * t_probe0.elapsed_time(t_probe1) * 1e-3) + 0.999))
    if world > 1:
        rr = torch.tensor([rounds], device=dev)
        dist.all_reduce(rr, op=dist.ReduceOp.MAX)
        rounds = int(rr.item())
    n_timed = args.steps * rounds
    phase_evs = [[ev() for _ in range(4)] for _ in range(n_timed)]
    e0, e1 = ev(), ev()
    barrier()
    e0.record()
    for k in range(n_timed):
        step(k, phase_evs[k])
    e1.record()
    barrier()
    total_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    ph = [sum(p[i].elapsed_time(p[i + 1]) for p in phase_evs) / n_timed for i in range(3)]
    if world > 1:
        tt = torch.tensor([total_ms] + ph, device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, ph = float(tt[0]), [float(x) for x in tt[1:]]
    roll_ms, batch_ms, tgt_ms = ph

    env_steps_per_step = N_GAMES * PLIES * world
    value = env_steps_per_step * n_timed / (total_ms * 1e-3)

    # ---------------- e2e: the same step through the public API with the replay index on the host ----------
    e2e = None
    if not args.no_e2e:
        # Software-pipelined like the reference's asynchronous batchers (train.py:261-309: selection and batch
        # making run beside the trainer).  The calling thread only ENQUEUES: rollout k, the D2H of its new episode
        # rows, the H2D of the windows of step k (drawn two steps earlier), make_batch, targets, the D2H of the
        # summary; it reads the summary of step k-2, which has long landed.  A background thread owns the replay
        # index: it waits for the rows of rollout k, commits them and draws the windows of step k+2
        # (b200rl_select_windows_host through ctypes: the GIL is released).  Nothing on the calling thread waits for
        # the GPU or for the selection unless the GPU queue is two steps deep, i.e. the run is GPU-bound.
        import queue
        import threading
        st = gen.store
        torch.cuda.synchronize()
        # slab of 65,536 rows per read-back: a rollout finishes ~13,400 episodes and commits lag up to two steps
        index = HostReplayIndex(gen, max_new_rows=1 << 16, n_threads=max(1, min(8, (os.cpu_count() or 1) // world)),
                                slots=4)
        DEPTH = 3                                            # window / summary slots in flight
        LOOK = 2                                             # windows of step k are drawn after rollout k-2
        win_pin = [torch.empty(B, 8, dtype=torch.int32).pin_memory() for _ in range(DEPTH)]
        win_devs = [torch.empty(B, 8, dtype=torch.int32, device=dev) for _ in range(DEPTH)]
        met_pin = [torch.zeros(4, dtype=torch.float32).pin_memory() for _ in range(DEPTH)]
        # PCIe traffic runs beside the kernels on its own streams, ordered by events.  Two streams: the D2H of a
        # rollout's rows has to wait for that rollout, the H2D of this step's windows must not queue behind it
        # (on one copy stream it did, and the GPU idled 0.16 ms per step waiting for its windows).  The summary of
        # the targets (4 means + D2H) runs on a third stream beside the next rollout.
        copy_stream = torch.cuda.Stream(device=dev)      # D2H: counters + finished-episode rows
        h2d_stream = torch.cuda.Stream(device=dev)       # H2D: windows
        sum_stream = torch.cuda.Stream(device=dev)       # target summary + its D2H
        ev_tail = [torch.cuda.Event() for _ in range(DEPTH)]
        ev_roll = [torch.cuda.Event() for _ in range(DEPTH)]
        ev_h2d = [torch.cuda.Event() for _ in range(DEPTH)]
        ev_done = [torch.cuda.Event() for _ in range(DEPTH)]
        win_ready = [threading.Event() for _ in range(DEPTH)]
        jobs = queue.Queue()
        hs = {"metric": 0.0, "d2h": 0, "select_s": 0.0, "select_n": 0, "err": None}

        def selector():
            """background: rows of rollout j -> commit -> windows of step j + LOOK"""
            try:
                torch.cuda.set_device(dev)
                while True:
                    job = jobs.get()
                    if job is None:
                        return
                    j, it_after, ticket = job
                    index.commit(ticket)                                         # waits: rows of rollout j have landed
                    t_s = pc()
                    slot = (j + LOOK) % DEPTH
                    index.select(B, TRAIN_ARGS, win_pin[slot], seed=rank, lookahead=LOOK * PLIES, it_now=it_after)
                    hs["select_s"] += pc() - t_s
                    hs["select_n"] += 1
                    hs["d2h"] = index.d2h_bytes
                    win_ready[slot].set()
            except Exception as ex:                                              # surface in the main thread
                hs["err"] = ex
                for w_ in win_ready:
                    w_.set()

        def step_tail(k, win):
            state["batch"] = batch = make_batch_windows(gen.store, win, TRAIN_ARGS, out=state["batch"])
            v, vr, rho, c = tsets[k % NSETS]
            t1, t2 = compute_target_heads("UPGO", "VTRACE", [(v, batch["outcome"], None, 1.0),
                                                              (vr, batch["return"], batch["reward"], 0.8)], 0.7, rho, c)
            return t1, t2

        e2e_evs = []
        tr_acc = {"enqueue_rollout_fetch": 0.0, "wait_windows": 0.0, "enqueue_tail": 0.0, "wait_result_2_steps_late": 0.0}
        pc = time.perf_counter

        def e2e_step(k):
            """k counts from 0 over warm-up + timed steps; steps 0..LOOK-1 have no windows yet (rollout only)"""
            t_a = pc()
            s = k % DEPTH
            evk = [ev() for _ in range(6)]
            e2e_evs.append(evk)
            evk[0].record()
            gen.rollout(PLIES, waves=waves, out_action=act_w)                   # async
            evk[1].record()
            ev_roll[s].record()
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ev_roll[s])
                ticket = index.fetch_async()                                    # D2H: counters + new episode rows
            evk[2].record()
            win_ready[(k + LOOK) % DEPTH].clear()
            jobs.put((k, gen.store.it, ticket))
            t_b = pc()
            t_c = t_b
            if k >= LOOK:
                win_ready[s].wait()                                             # drawn two steps ago: normally set
                if hs["err"] is not None:
                    raise hs["err"]
                t_c = pc()
                with torch.cuda.stream(h2d_stream):                             # H2D of this step's windows
                    win_devs[s].copy_(win_pin[s], non_blocking=True)
                    ev_h2d[s].record()
                torch.cuda.current_stream().wait_event(ev_h2d[s])
                evk[3].record()
                (tv, av), (tr, ar) = step_tail(k, win_devs[s])
                evk[4].record()
                ev_tail[s].record()
                with torch.cuda.stream(sum_stream):                             # result of the step -> host
                    sum_stream.wait_event(ev_tail[s])
                    # the four outputs are views of ONE (2, 2, B, T, P, 1) buffer (compute_target_heads): one reduction
                    base = tv._base if (tv._base is not None and tv._base.numel() == 4 * tv.numel()) else \
                        torch.stack([tv, av, tr, ar])
                    base.record_stream(sum_stream)
                    met_pin[s].copy_(base.view(4, -1).mean(dim=1), non_blocking=True)
                    ev_done[s].record()
                evk[5].record()
            t_d = pc()
            if k >= 2 * LOOK:                                                   # result of step k-2 (D2H landed long ago)
                ev_done[(k - LOOK) % DEPTH].synchronize()
                hs["metric"] = float(met_pin[(k - LOOK) % DEPTH][0])
            t_e = pc()
            for name, dt_ in zip(tr_acc, (t_b - t_a, t_c - t_b, t_d - t_c, t_e - t_d)):
                tr_acc[name] += dt_

        th = threading.Thread(target=selector, daemon=True)
        th.start()
        n_warm = 6
        for k in range(n_warm):
            e2e_step(k)
        barrier()
        ksteps = max(5, min(args.steps, 20))
        e2e_rounds = max(1, int(0.35 / (ksteps * total_ms / (args.steps * rounds) * 1e-3) + 0.999))
        for name in tr_acc:
            tr_acc[name] = 0.0
        hs["select_s"], hs["select_n"] = 0.0, 0
        e2e_evs.clear()
        t0 = time.perf_counter()
        for k in range(ksteps * e2e_rounds):
            e2e_step(n_warm + k)
        last = n_warm + ksteps * e2e_rounds - 1
        for k in (last - 1, last):                                              # drain the two results still in flight
            ev_done[k % DEPTH].synchronize()
            hs["metric"] = float(met_pin[k % DEPTH][0])
        barrier()
        dt = time.perf_counter() - t0
        jobs.put(None)
        th.join(timeout=10)
        n_e2e = ksteps * e2e_rounds
        if world > 1:
            tt = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        timed_evs = [e_ for e_ in e2e_evs if True]
        e2e = {"value": N_GAMES * PLIES * n_e2e * world / dt, "unit": "env-steps/s",
               "h2d_bytes_per_step": int(win_devs[0].numel() * 4),
               "d2h_bytes_per_step": int(hs["d2h"] + met_pin[0].numel() * 4),
               "steps": n_e2e, "ms_per_step": dt / n_e2e * 1e3,
               "host_ms_per_step": {name: round(v_ / n_e2e * 1e3, 4) for name, v_ in tr_acc.items()},
               "selector_thread_ms_per_step": round(hs["select_s"] / max(hs["select_n"], 1) * 1e3, 4),
               "selector_threads": index.n_threads,
               "gpu_ms_per_step": {name: round(sum(e_[i].elapsed_time(e_[i + 1]) for e_ in timed_evs) / len(timed_evs), 4)
                                   for i, name in enumerate(("rollout", "enqueue_gap", "wait_h2d_windows",
                                                             "make_batch_targets", "gap_after_targets"))},
               "note": "per step, calling thread (enqueue only): BatchGenerator.rollout; D2H of the counters and the "
                       "new finished-episode rows (copy stream); pinned H2D of the windows drawn two steps earlier (second copy "
                       "stream, event-ordered); make_batch_windows; compute_target_heads (value + return head, one launch); "
                       "target summary + its D2H on a third stream, read two steps late.  Background thread: HostReplayIndex.commit + "
                       "select (b200rl_select_windows_host, replay index on the host).  Moments stay in HBM (the "
                       "trajectory ring is the replay buffer).  Wall clock incl. all host work."}

        # the same self-play driven ply by ply from HOST buffers (reference Environment-style stepping: actions
        # H2D, observation / legal bits / reward / terminal D2H every ply) — PCIe-bound by the fp32 observations
        from handyrl_b200.engine import BatchedEnv
        henv = BatchedEnv("Geister", N_GAMES, device=dev)
        out = henv.step_observe(None)
        keys = ("main", "scalar", "action_mask", "reward", "terminal")
        h = {k2: torch.empty(out[k2].shape, dtype=out[k2].dtype).pin_memory() for k2 in keys}
        h_act = torch.zeros(N_GAMES, dtype=torch.int32).pin_memory()
        h_idx = torch.zeros(N_GAMES, dtype=torch.int64)
        d_act = torch.empty(N_GAMES, dtype=torch.int32, device=dev)
        for k2 in keys:
            h[k2].copy_(out[k2])
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(world, 1)))

        def host_ply():
            nonlocal out
            torch.argmin(h["action_mask"], dim=1, out=h_idx)                # host policy: first legal action
            h_act.copy_(h_idx)
            d_act.copy_(h_act, non_blocking=True)                           # H2D
            out = henv.step_observe(d_act, auto_reset=True, out=out)
            for k2 in keys:
                h[k2].copy_(out[k2], non_blocking=True)                     # D2H
            torch.cuda.current_stream().synchronize()

        for _ in range(4):
            host_ply()
        barrier()
        nply = 16
        t0 = time.perf_counter()
        for _ in range(nply):
            host_ply()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        d2h = sum(t_.numel() * t_.element_size() for t_ in h.values())
        e2e["host_stepping"] = {"value": N_GAMES * nply * world / dt, "unit": "env-steps/s",
                                "h2d_bytes_per_ply": int(h_act.numel() * 4), "d2h_bytes_per_ply": int(d2h),
                                "pcie_gbs": d2h * nply / dt / 1e9,
                                "note": "BatchedEnv.step_observe per ply with pinned-host actions in and fp32 "
                                        "observation/action_mask/reward/terminal out (1.96 kB per env-step over PCIe)"}

    learner = None
    if not args.no_learner:
        learner = learner_leg(dev, rank, world, barrier)
        if world > 1:
            # the reference's workers with the real net (batch-of-1 CPU forward per ply, worker.py:52, model.py:43-53)
            # on this box's cores, timed by rank 0 while the other ranks wait
            if rank == 0:
                ref = cpu_baseline_reference(5.0, parts="generation_net")
                learner["reference_workers"] = {
                    "env_steps_per_s": (ref.get("rates") or {}).get("generation_real_net_env_steps_per_s"),
                    "cores": ref.get("cores"), "kind": "reference",
                    "sample": "Generator.generate, Geister, ModelWrapper(GeisterNet), os.cpu_count() processes, 5 s"}
            dist.barrier()

    if rank == 0:
        peak, peak_kind = load_peaks()

        def roof(kernel, nbytes, ms, unit_bytes, unit_name):
            a = nbytes / (ms * 1e-3) / 1e9
            return {"kernel": kernel, "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                    "traffic": None, "peak_kind": peak_kind, "bytes_per_unit": unit_bytes, "unit_name": unit_name,
                    "avg_ms": ms}

        roofs = {
            "rollout": roof("k_rollout<Geister>", BYTES_PER_ENV_STEP * N_GAMES * PLIES, roll_ms, BYTES_PER_ENV_STEP,
                            "env-step"),
            "make_batch": roof("k_select_windows + k_make_batch<Geister>", BYTES_PER_BATCH_CELL * B * T, batch_ms,
                               BYTES_PER_BATCH_CELL, "(b,t) batch cell"),
            "targets": roof("k_target_rows2<UPGO,VTRACE> (value + return head, one launch)", BYTES_PER_TARGET_PAIR * B * T * P,
                            tgt_ms, BYTES_PER_TARGET_PAIR, "(b,t,p) element, both heads"),
        }
        dominant = max(roofs, key=lambda k: roofs[k]["avg_ms"])
        # dram bytes per launch from the ncu --set full capture of this workload (tools/measure_round.sh ->
        # tools/make_traffic.py).  Only valid for the sources it was captured with: a stale file reads as null.
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file):
            with open(traffic_file) as f:
                tr = json.load(f)
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            from make_traffic import source_hash
            fresh = tr.get("_source_hash") == source_hash()
            for k in roofs:
                roofs[k]["traffic"] = tr.get(k) if fresh else None
                roofs[k]["traffic_source"] = tr.get("_source") if fresh else (
                    "stale: profiles/traffic.json was captured with other CUDA sources (re-run tools/measure_round.sh)")
        line = {
            "metric": "geister_selfplay_env_steps_per_sec", "value": value, "unit": "env-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / n_timed,
            "rounds": rounds, "timed_steps": n_timed, "timed_region_ms": total_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64 bitboards -> f32", "data": "synthetic", "config": workload_config(world),
            "target_timesteps_per_sec": B * T * P * world / (tgt_ms * 1e-3),
            "batch_timesteps_per_sec": B * T * world / (batch_ms * 1e-3),
            "phases": {"rollout_ms": roll_ms, "make_batch_ms": batch_ms, "targets_ms": tgt_ms,
                       "env_steps_per_sec_rollout_only": env_steps_per_step / (roll_ms * 1e-3)},
            "roofline": roofs[dominant], "rooflines": roofs,
            "clocks": clocks, "e2e": e2e, "gpu_launches": 4 * n_timed,
            "extra": {"learner": learner},
        }
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline()
            line["cpu_baseline_reference"] = cpu_baseline_reference()
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
