"""compute_target sweep (BASELINE.json configs[4]): T in {32,200,1024} x B in {4096,65536} x P=2,
all four modes + the fused V-Trace/UPGO pair, return head (rewards given) and value head (rewards None).
Prints one JSON object; run on the GPU box:  python tools/sweep_targets.py > gpurun_out/sweep.json
Under torchrun (N ranks, one per GPU) every rank runs the sweep on its own shard - SWEEP_SCALING=weak (default): B rows
per GPU, whole-job rate = N x; strong: B / N rows per GPU (E1: target rows shard over the batch dimension, no
collective) - the per-launch time is the MAX over the ranks; rank 0 prints.  SWEEP_CPU=1 (rank 0) adds the reference's
losses.compute_target on this box's CPU beside each cell (tools/time_reference.py --sweep-targets).
B200RL_TARGET_GENERIC=1 forces the generic strided kernel (A/B against the row-per-thread fast path).
SWEEP_ALIAS=1 passes ONE tensor as rhos and cs (what learner.compute_loss does): 2 B/element less for V-Trace."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200.losses import compute_target, compute_target_pair  # noqa: E402

PEAK = 6548.8
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
# algorithmic bytes per (b,t,p) element at P=2: with rewards / without
BYTES = {"MC": (12, 8), "TD": (16, 12), "UPGO": (16, 12), "VTRACE": (20, 16), "VTRACE+UPGO": (20, 16)}


ALIAS = bool(os.environ.get("SWEEP_ALIAS"))


def bench(fn, sets, iters, graph=False):
    """ms per call.  graph=True: the `iters` calls are captured into one CUDA graph and replayed, so that a
    kernel shorter than the ~18 us host cost of a call is timed on the device and not by the launch rate."""
    for i in range(3):
        fn(sets[i % len(sets)])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(iters):
                fn(sets[i % len(sets)])
        g.replay()
        torch.cuda.synchronize()
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        del g
        return ms
    e0.record()
    for i in range(iters):
        fn(sets[i % len(sets)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    import torch.distributed as dist
    sys.stdout.flush()
    result_out = os.fdopen(os.dup(1), "w")      # stdout carries only the JSON (NCCL prints its banner to fd 1)
    os.dup2(2, 1)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    scaling = os.environ.get("SWEEP_SCALING", "weak")
    g = torch.Generator(device=dev).manual_seed(rank)
    out = {"peak_gbs": PEAK, "generic": bool(os.environ.get("B200RL_TARGET_GENERIC")), "cs_is_rhos": ALIAS,
           "n_gpus": world, "scaling": scaling, "rows": []}
    Ts = [int(x) for x in os.environ.get('SWEEP_T', '32,200,1024').split(',')]
    Bs = [int(x) for x in os.environ.get('SWEEP_B', '4096,65536').split(',')]
    for T in Ts:
        for B_total in Bs:
            P = 2
            B = B_total if scaling == "weak" else max(32, B_total // world)
            n = B * T * P
            nsets = max(2, min(8, int(300e6 // (n * 4 * 4)) + 1))   # rotate inputs past the 126 MB L2
            sets = []
            for _ in range(nsets):
                v = torch.tanh(torch.randn(B, T, P, 1, device=dev, generator=g))
                rew = torch.full((B, T, P, 1), -0.01, device=dev)
                ret = torch.randn(B, T, P, 1, device=dev, generator=g)
                oc = torch.randint(-1, 2, (B, 1, P, 1), device=dev, generator=g).float()
                rho = torch.exp(0.5 * torch.randn(B, T, 1, 1, device=dev, generator=g)).clamp(0, 1)
                sets.append((v, rew, ret, oc, rho, rho if ALIAS else rho.clone()))
            iters = 20 if n > 5e7 else 50
            for head in ("return", "value"):
                for mode in ("MC", "TD", "UPGO", "VTRACE", "VTRACE+UPGO"):
                    if mode == "VTRACE+UPGO":
                        if head == "return":
                            fn = lambda s: compute_target_pair("VTRACE", "UPGO", s[0], s[2], s[1], 0.7, 0.8, s[4], s[5])
                        else:
                            fn = lambda s: compute_target_pair("VTRACE", "UPGO", s[0], s[3], None, 0.7, 1.0, s[4], s[5])
                    elif head == "return":
                        fn = lambda s, m=mode: compute_target(m, s[0], s[2], s[1], 0.7, 0.8, s[4], s[5])
                    else:
                        fn = lambda s, m=mode: compute_target(m, s[0], s[3], None, 0.7, 1.0, s[4], s[5])
                    graph = n <= (1 << 24)      # short kernels: replay a captured graph
                    ms = bench(fn, sets, iters, graph)
                    # kernels shorter than a Python call: also the rate a caller sees issuing the calls one by one
                    ms_call = bench(fn, sets, iters, False) if graph else ms
                    if world > 1:
                        tt = torch.tensor([ms], device=dev, dtype=torch.float64)
                        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                        ms = float(tt.item())
                    by = BYTES[mode][0 if head == "return" else 1] - (2 if ALIAS and "VTRACE" in mode else 0)
                    gbs = n * by / (ms * 1e-3) / 1e9
                    out["rows"].append({"T": T, "B": B_total, "B_per_gpu": B, "P": P, "head": head, "mode": mode,
                                        "ms": round(ms, 4), "elements_per_s": n * world / (ms * 1e-3),
                                        "elements_per_s_per_gpu": n / (ms * 1e-3), "bytes_per_element": by, "timed": "graph" if graph else "stream",
                                        "ms_per_python_call": round(ms_call, 4),
                                        "GBps": round(gbs, 1), "frac_of_peak": round(gbs / PEAK, 3)})
            del sets
            torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if os.environ.get("SWEEP_CPU"):
        import subprocess
        try:
            r = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                             "time_reference.py"), "--sweep-targets"],
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=900, text=True)
            cpu = json.loads(r.stdout.strip().splitlines()[-1])
            key = {(c["T"], c["B"], c["head"], c["mode"]): c for c in cpu["rows"]}
            for row in out["rows"]:
                c = key.get((row["T"], row["B"], row["head"], row["mode"]))
                if c:
                    row["cpu_reference_elements_per_s_1_thread"] = c["elements_per_s_1_thread"]
                    row["cpu_reference_elements_per_s_all_threads"] = c["elements_per_s_all_threads"]
            out["cpu_reference"] = {k: cpu[k] for k in ("cores", "kind", "sample")}
        except Exception as ex:
            out["cpu_reference"] = {"unavailable": repr(ex)}
    print(json.dumps(out), file=result_out, flush=True)
    if world > 1:
        dist.destroy_process_group()
    for r in out["rows"]:
        print("T=%4d B=%5d %-6s %-12s %8.3f ms %7.1f Gelem/s %7.1f GB/s  %.3f" % (
            r["T"], r["B"], r["head"], r["mode"], r["ms"], r["elements_per_s"] / 1e9, r["GBps"], r["frac_of_peak"]),
            file=sys.stderr)


if __name__ == "__main__":
    main()
