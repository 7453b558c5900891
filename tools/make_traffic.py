"""profiles/traffic.json from one `ncu --set full` capture of `bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline
--no-learner` (tools/measure_round.sh): DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per launch of the
kernels of each bench phase, plus the hash of the CUDA sources the capture was taken with.  bench.py reports
`roofline.traffic` from this file ONLY while that hash matches the sources it runs (a stale capture reads as null), and
this script fails if a phase has no kernel in the capture (the kernel list changed).

    python tools/make_traffic.py gpurun_out/r2/step_full.ncu-rep profiles/traffic.json"""
import csv
import glob
import hashlib
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "Tbyte": 1e12}
PHASES = {"rollout": ("k_rollout",), "make_batch": ("k_select_windows", "k_make_batch"),
          "targets": ("k_target_rows2", "k_target_rows", "k_target_scan")}


def source_hash():
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "handyrl_b200", "csrc", "*.cu*"))):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    per = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        b = sum(float(r[col[k]]) * UNIT[units[col[k]]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        per.setdefault(name, []).append(b)
    res = {"_source": "%s (ncu --set full --clock-control none, per launch, bench.py default workload)" % os.path.basename(rep),
           "_source_hash": source_hash(), "_kernels": {}}
    for phase, keys in PHASES.items():
        tot, found = 0.0, []
        for name, v in per.items():
            # ncu prints e.g. "void <unnamed>::k_rollout<GeisterGame<0>, 0>(<unnamed>::RolloutParams)"
            m = re.search(r"\b(k_[A-Za-z0-9_]+)", name)
            if m and m.group(1) in keys:
                tot += sum(v) / len(v)
                found.append(name[:100])
        if not found:
            sys.exit("make_traffic: no kernel of phase '%s' (%s) in %s - the kernel list changed" % (phase, keys, rep))
        res[phase] = int(tot)
        res["_kernels"][phase] = found
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
