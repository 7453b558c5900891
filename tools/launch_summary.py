"""Summarise the ncu launch list of tools/measure_round.sh (`--metrics gpu__time_duration.sum`) as markdown: launches,
mean time and share per kernel, and the per-STEP shares next to the CUDA-event phases of the un-profiled bench line.
    python tools/launch_summary.py gpurun_out/r2b/launches.csv gpurun_out/r2b/bench.json > profiles/r2_launches.md"""
import collections
import csv
import json
import sys

UNIT = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}


def main(launches, bench):
    rows = [r for r in csv.reader(l for l in open(launches) if l.startswith('"'))]
    col = {h: i for i, h in enumerate(rows[0])}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if r[col["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v = float(r[col["Metric Value"]].replace(",", "")) * UNIT.get(r[col["Metric Unit"]], 1e-3)
        agg.setdefault(r[col["Kernel Name"]], []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print("# ncu launch list of `bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-learner --rounds 1`\n")
    print("`ncu --metrics gpu__time_duration.sum --clock-control none -c 400` (tools/measure_round.sh, fresh 1xB200). Per-launch times")
    print("under ncu are cold-cache and serialised: what must agree with the CUDA-event phases of bench.py is each kernel's SHARE.\n")
    print("| kernel | launches | mean us | share of all captured time |\n|---|---|---|---|")
    for n, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("| `%s` | %d | %.1f | %.1f %% |" % (n[:110].replace("|", "/"), len(v), sum(v) / len(v), 100 * sum(v) / tot))

    def mean(key):
        v = [x for n, vs in agg.items() if key in n for x in vs]
        return sum(v) / len(v) if v else 0.0
    ro, sw, mb, tg = mean("k_rollout"), mean("k_select_windows"), mean("k_make_batch"), mean("k_target_rows")
    step = ro + sw + mb + tg
    ph = json.load(open(bench))["phases"]
    ev = ph["rollout_ms"] + ph["make_batch_ms"] + ph["targets_ms"]
    print("\nThe capture also holds the untimed prefill rollouts and warm-up steps, so the last column over-weights the rollout.")
    print("Per STEP (one launch of each, mean times above): rollout %.1f us = %.1f %%, k_select_windows + k_make_batch %.1f us = "
          "%.1f %%, k_target_rows2 %.1f us = %.1f %% - against %.1f / %.1f / %.1f %% from CUDA events in the un-profiled run (the event "
          "bracket of the targets phase also holds a launch gap)." % (
              ro, 100 * ro / step, sw + mb, 100 * (sw + mb) / step, tg, 100 * tg / step,
              100 * ph["rollout_ms"] / ev, 100 * ph["make_batch_ms"] / ev, 100 * ph["targets_ms"] / ev))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
