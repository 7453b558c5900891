"""Attribute the executed SASS instructions of one kernel (ncu source page, CSV) to CUDA-C source lines.

    ncu -i cap.ncu-rep --page source --csv > cap_source.csv          (on the GPU box)
    python tools/sass_lines.py cap_source.csv env k_rolloutI13TicTacToeGameILb0EELb0EE 524288 [--inline-of env.cu:418]

The ncu CSV lists one row per SASS instruction in address order; the same cubin, extracted from the library that ran
(built with -lineinfo), is disassembled here with `nvdisasm -gi`, whose `//## File ..., line N [inlined at ...]`
comments give the innermost source line of every instruction.  Output: warp instructions per UNIT (pass the number of
units the launch processed, e.g. warp-plies) by source line, with the opcode mix and the stall samples of each line.
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def disassemble(cubin_name, func_pat):
    tmp = tempfile.mkdtemp()
    lib = os.path.join(ROOT, "handyrl_b200", "libb200rl.so")
    subprocess.run(["cuobjdump", "-xelf", cubin_name + ".sm_100a.cubin", lib], cwd=tmp, check=True,
                   stdout=subprocess.DEVNULL)
    txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cubin_name + ".sm_100a.cubin")],
                         capture_output=True, text=True).stdout.splitlines()
    start = [i for i, l in enumerate(txt) if l.startswith("//-") and ".text." in l]
    pick = [i for i in start if func_pat in txt[i]]
    if len(pick) != 1:
        sys.exit("function pattern matches %d sections" % len(pick))
    end = min([i for i in start if i > pick[0]] + [len(txt)])
    line_of, group, tagged = {}, [], False
    for l in txt[pick[0]:end]:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            if not tagged:
                group = []
            group.append((os.path.basename(m.group(1)), int(m.group(2))))
            tagged = True
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/\s+\S', l)
        if m:
            line_of[int(m.group(1), 16)] = tuple(group)
            tagged = False
    return line_of


def main():
    src_csv, cubin, func, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
    only = None
    if "--inline-of" in sys.argv:
        f, n = sys.argv[sys.argv.index("--inline-of") + 1].split(":")
        only = (f, int(n))
    rows = list(csv.reader(open(src_csv)))
    hdr, data = rows[1], rows[2:]
    ia, ie, isrc, iss = (hdr.index(k) for k in ("Address", "Instructions Executed", "Source", "# Samples"))
    base = int(data[0][ia], 16)
    line_of = disassemble(cubin, func)
    by, smp, ops = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
    for r in data:
        g = line_of.get(int(r[ia], 16) - base, ())
        if only and only not in g:
            continue
        k = g[0] if g else None
        s = r[isrc].split()
        op = s[1] if s[0].startswith("@") else s[0]
        by[k] += int(r[ie])
        smp[k] += int(r[iss])
        ops[k][op.split(".")[0]] += int(r[ie])
    print("warp instructions per unit: %.1f   (stall samples %d)" % (sum(by.values()) / units, sum(smp.values())))
    for k, v in by.most_common(50):
        print("%-32s %7.1f  samples %6d  %s" % ("%s:%d" % k if k else "?", v / units, smp[k],
                                               " ".join("%s:%.1f" % (o, c / units) for o, c in ops[k].most_common(6))))
    tot = collections.Counter()
    for k in ops:
        tot.update(ops[k])
    print("opcode mix per unit: " + " ".join("%s:%.1f" % (o, c / units) for o, c in tot.most_common(24)))


if __name__ == "__main__":
    main()
