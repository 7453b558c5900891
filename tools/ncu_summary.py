"""Summarise an .ncu-rep (one `ncu --set full` capture) as markdown + the per-launch DRAM traffic.
    python tools/ncu_summary.py gpurun_out/r1_step_full.ncu-rep > profiles/r1_step_full.md"""
import csv
import io
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram read"),
        ("dram__bytes_write.sum", "dram write"), ("dram__bytes.sum.per_second", "dram rate"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of ncu peak"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/TEX %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("smsp__inst_executed.sum", "warp instructions"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs/thread"),
        ("launch__shared_mem_per_block_static", "static smem"), ("launch__shared_mem_per_block_dynamic", "dynamic smem"),
        ("launch__waves_per_multiprocessor", "waves/SM"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts")]
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "Tbyte": 1e12}


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    traffic = {}
    print("# ncu --set full --clock-control none: %s\n" % path.split("/")[-1])
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        print("## `%s`\n" % name[:110])
        print("| metric | value |\n|---|---|")
        for k, label in KEYS:
            if k in col:
                print("| %s (`%s`) | %s %s |" % (label, k, r[col[k]], units[col[k]]))
        st = [(h, float(r[i])) for h, i in col.items() if h.startswith("smsp__pcsamp_warps_issue_stalled_")
              and not h.endswith("_not_issued") and r[i]]
        tot = sum(v for _, v in st) or 1
        top = sorted(st, key=lambda x: -x[1])[:6]
        print("| top stall samples | %s |" % ", ".join("%s %.0f%%" % (h.replace("smsp__pcsamp_warps_issue_stalled_", ""),
                                                                     100 * v / tot) for h, v in top))
        try:
            b = sum(float(r[col[k]]) * UNIT[units[col[k]]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            traffic.setdefault(name.split("(")[0].split("::")[-1], []).append(b)
            print("| **traffic per launch** | %.4g GB |" % (b / 1e9))
        except Exception:
            pass
        print()
    print("<!-- traffic: %r -->" % {k: sum(v) / len(v) for k, v in traffic.items()})


if __name__ == "__main__":
    main(sys.argv[1])
