"""Time the UNMODIFIED reference's Python hot path on this machine's host cores.

The reference tree is the git-ignored install `baseline/_ref/` (tools/install_ref.sh).  `procs` OS processes (default
os.cpu_count(), the reference's `num_parallel` / `num_batchers` model: one single-threaded process per core,
main.py:10, model.py:8) each run, for a fixed wall time per part:

  generation   handyrl.generation.Generator.generate on the reference's Geister env with ModelWrapper(RandomModel)
               (the epoch-0 policy, train.py:438-439; includes the bz2+pickle of generation.py:79-86)
  generation_net  the same with the reference's GeisterNet (batch-of-1 CPU forward per ply, model.py:43-53)
  make_batch   handyrl.train.make_batch on windows drawn by Batcher.select_episode from those episodes
               (batch_size 256, forward_steps T; includes the bz2 decode of train.py:54)
  targets      handyrl.losses.compute_target as compute_loss calls it (train.py:245-253): value_target UPGO and
               policy_target VTRACE on the value head (bootstrap = outcome, no rewards, gamma 1) and the return head
               (rewards, gamma 0.8) of a (B, T, 2, 1) batch = 4 calls

and the rates are summed over the processes.  `whole_step` combines them into the metric of bench.py (env-steps/s of
generate -> make_batch -> targets where every generated ply is batched and targeted once).

This is measurement infrastructure for bench.py's `cpu_baseline_reference` object; nothing here is on a product path.

    python tools/time_reference.py --seconds 6 [--procs N] [--parts generation,make_batch,targets]
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")

ARGS = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 32, 'compress_steps': 4,
        'maximum_episodes': 100000, 'batch_size': 256, 'lambda': 0.7}


def _worker(job):
    wid, parts, seconds, env_name, T, tB = job
    sys.path.insert(0, REF)
    import random
    from types import SimpleNamespace
    import numpy as np
    import torch
    from handyrl.environment import make_env
    from handyrl.generation import Generator
    from handyrl.model import ModelWrapper, RandomModel            # sets torch.set_num_threads(1) (model.py:8)
    from handyrl.train import Batcher, make_batch
    from handyrl.losses import compute_target
    random.seed(1000 + wid)
    torch.manual_seed(wid)
    args = dict(ARGS, forward_steps=T)
    env = make_env({'env': env_name, 'id': wid})
    gen = Generator(env, args)
    out = {}
    episodes = []

    def run_generation(models, secs, keep):
        steps = n = 0
        gen.generate(models, {})                                   # warm-up, untimed
        t0 = time.perf_counter()
        while True:
            ep = gen.generate(models, {})
            steps += ep['steps']
            n += 1
            if keep and len(episodes) < 64:
                episodes.append(ep)
            dt = time.perf_counter() - t0
            if dt >= secs:
                return steps, n, dt

    if 'generation' in parts or 'make_batch' in parts:
        models = {p: ModelWrapper(RandomModel(env)) for p in env.players()}
        s, n, dt = run_generation(models, seconds, True)
        out['generation'] = (s, n, dt)
    if 'generation_net' in parts:
        net = env.net()()
        net.eval()
        models = {p: ModelWrapper(net) for p in env.players()}
        out['generation_net'] = run_generation(models, seconds, False)
    if 'make_batch' in parts:
        stub = SimpleNamespace(episodes=episodes, args=args)
        cells = calls = 0
        make_batch([Batcher.select_episode(stub) for _ in range(8)], args)     # warm-up
        t0 = time.perf_counter()
        while True:
            make_batch([Batcher.select_episode(stub) for _ in range(args['batch_size'])], args)
            cells += args['batch_size'] * T
            calls += 1
            dt = time.perf_counter() - t0
            if dt >= seconds:
                break
        out['make_batch'] = (cells, calls, dt)
    if 'targets' in parts:
        g = torch.Generator().manual_seed(wid)
        P = 2
        v = torch.tanh(torch.randn(tB, T, P, 1, generator=g))
        vr = torch.tanh(torch.randn(tB, T, P, 1, generator=g))
        outcome = torch.randint(-1, 2, (tB, 1, P, 1), generator=g).float()
        rew = torch.full((tB, T, P, 1), -0.01)
        ret = torch.zeros(tB, T, P, 1)
        acc = torch.zeros(tB, P, 1)
        for t in range(T - 1, -1, -1):
            acc = rew[:, t] + 0.8 * acc
            ret[:, t] = acc
        rho = torch.exp(0.5 * torch.randn(tB, T, 1, 1, generator=g)).clamp(0, 1)
        c = rho.clone()

        def once():
            compute_target('UPGO', v, outcome, None, 0.7, 1, rho, c)
            compute_target('UPGO', vr, ret, rew, 0.7, 0.8, rho, c)
            compute_target('VTRACE', v, outcome, None, 0.7, 1, rho, c)
            compute_target('VTRACE', vr, ret, rew, 0.7, 0.8, rho, c)
        once()
        n = 0
        t0 = time.perf_counter()
        while True:
            once()
            n += 1
            dt = time.perf_counter() - t0
            if dt >= seconds:
                break
        out['targets'] = (n * tB * T * P, n, dt)
    return out


def measure(seconds=6.0, procs=None, parts=('generation', 'make_batch', 'targets'), env_name='Geister', T=32,
            target_rows=1024):
    procs = procs or os.cpu_count() or 1
    if not os.path.exists(os.path.join(REF, "handyrl", "train.py")):
        return {"unavailable": "baseline/_ref not installed (tools/install_ref.sh needs /root/reference)"}
    os.environ["OMP_NUM_THREADS"] = "1"                              # main.py:10
    sys.path.insert(0, REF)
    import torch                                                     # noqa: F401  imported once, shared by the forks
    import handyrl.train                                             # noqa: F401
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(procs) as pool:
        res = pool.map(_worker, [(w, tuple(parts), seconds, env_name, T, target_rows) for w in range(procs)])
    wall = time.perf_counter() - t0
    rate = {}
    detail = {}
    for part in ('generation', 'generation_net', 'make_batch', 'targets'):
        rows = [r[part] for r in res if part in r]
        if rows:
            rate[part] = sum(a / dt for a, _, dt in rows)            # units/s summed over the processes
            detail[part] = {"units": int(sum(a for a, _, _ in rows)), "calls": int(sum(n for _, n, _ in rows)),
                            "seconds_per_process": round(sum(dt for _, _, dt in rows) / len(rows), 3)}
    out = {"cores": procs, "kind": "reference", "wall_seconds": round(wall, 1),
           "rates": {"generation_env_steps_per_s": rate.get('generation'),
                     "generation_real_net_env_steps_per_s": rate.get('generation_net'),
                     "make_batch_timesteps_per_s": rate.get('make_batch'),
                     "target_elements_per_s_both_heads_both_modes": rate.get('targets')},
           "detail": detail}
    if all(k in rate for k in ('generation', 'make_batch', 'targets')):
        # one env-step of the whole step = 1 generated ply + 1 batch cell + 2 players' target elements
        per_step = 1.0 / rate['generation'] + 1.0 / rate['make_batch'] + 2.0 / rate['targets']
        out["value"] = 1.0 / per_step
        out["unit"] = "env-steps/s"
    out["sample"] = ("UNMODIFIED reference (baseline/_ref) in %d single-threaded processes, %.0f s per part: "
                     "Generator.generate Geister + ModelWrapper(RandomModel) incl. bz2/pickle; make_batch of 256 "
                     "windows x T=%d drawn by Batcher.select_episode incl. bz2 decode; losses.compute_target UPGO + "
                     "VTRACE on the value and return heads of (%d,%d,2,1); whole-step value = 1/(1/gen + 1/batch + "
                     "2/targets)" % (procs, seconds, T, target_rows, T))
    return out


def sweep_targets_cpu(cells=((32, 4096), (200, 4096), (1024, 4096), (32, 65536)), budget=0.4):
    """BASELINE configs[4] on the host: the reference's losses.compute_target on CPU tensors of tools/sweep_targets.py's
    cells, with 1 torch thread (what the reference runs with: main.py:10, model.py:8) and with all cores.  Element
    rates per (T, B, head, mode); B = 65,536 only at T = 32."""
    sys.path.insert(0, REF)
    import torch
    from handyrl.losses import compute_target
    cores = os.cpu_count() or 1
    rows = []
    g = torch.Generator().manual_seed(0)
    for T, B in cells:
        P = 2
        v = torch.tanh(torch.randn(B, T, P, 1, generator=g))
        rew = torch.full((B, T, P, 1), -0.01)
        ret = torch.randn(B, T, P, 1, generator=g)
        oc = torch.randint(-1, 2, (B, 1, P, 1), generator=g).float()
        rho = torch.exp(0.5 * torch.randn(B, T, 1, 1, generator=g)).clamp(0, 1)
        c = rho.clone()
        n = B * T * P
        for head in ("return", "value"):
            for mode in ("MC", "TD", "UPGO", "VTRACE", "VTRACE+UPGO"):
                def call():
                    for m in (mode.split("+")):
                        if head == "return":
                            compute_target(m, v, ret, rew, 0.7, 0.8, rho, c)
                        else:
                            compute_target(m, v, oc, None, 0.7, 1.0, rho, c)
                rates = {}
                for threads in (1, cores):
                    torch.set_num_threads(threads)
                    call()
                    call()                                  # the first call after a thread-count change builds the pool
                    k, t0 = 0, time.perf_counter()
                    while True:
                        call()
                        k += 1
                        dt = time.perf_counter() - t0
                        if dt >= budget or k >= 50:
                            break
                    rates[threads] = n * k / dt
                rows.append({"T": T, "B": B, "head": head, "mode": mode, "elements_per_s_1_thread": rates[1],
                             "elements_per_s_all_threads": rates[cores]})
    return {"cores": cores, "kind": "reference", "rows": rows,
            "sample": "handyrl.losses.compute_target (baseline/_ref) on CPU tensors, >= %.1f s or 50 calls per cell; "
                      "VTRACE+UPGO = the two calls of train.py:248-253" % budget}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=6.0)
    ap.add_argument("--procs", type=int, default=0)
    ap.add_argument("--parts", default="generation,make_batch,targets")
    ap.add_argument("--env", default="Geister")
    ap.add_argument("--T", type=int, default=32)
    ap.add_argument("--target-rows", type=int, default=1024)
    ap.add_argument("--sweep-targets", action="store_true")
    a = ap.parse_args()
    if a.sweep_targets:
        print(json.dumps(sweep_targets_cpu()))
        sys.exit(0)
    print(json.dumps(measure(a.seconds, a.procs or None, tuple(a.parts.split(',')), a.env, a.T, a.target_rows)))
