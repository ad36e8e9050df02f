#!/usr/bin/env python
"""Per-source-line instruction and stall-sample totals for one kernel of an ncu report.

ncu's CSV source page is SASS only; this joins it, instruction by instruction, with `nvdisasm -g` line markers of
the in-tree library (both list the kernel's instructions in address order).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep k_fast [top]
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "snake-slam_b200", "libsnakeorb.so")


def sass_lines(kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, capture_output=True)
    for f in sorted(os.listdir(tmp)):
        txt = subprocess.run(["nvdisasm", "-g", "-c", f], cwd=tmp, capture_output=True, text=True).stdout
        out, on, line = [], False, 0
        for l in txt.splitlines():
            if l.startswith("\t.section"):
                on = (".text." in l) and (kernel in l)
                continue
            if not on:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if m:
                line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
                out.append(line)
        if out:
            return out
    raise SystemExit("kernel not found in " + LIB)


def main(rep, kernel, top=40):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kernel], check=True,
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    head = rows[h]
    ci, cs = head.index("Instructions Executed"), head.index("# Samples")
    body = []
    for r in rows[h + 1:]:
        if r and r[0] == "Kernel Name":
            break  # first launch only
        if len(r) == len(head):
            body.append(r)
    lines = sass_lines(kernel)
    if len(lines) != len(body):
        print("warning: %d SASS instructions in the library, %d in the report" % (len(lines), len(body)))
    inst, samp = collections.Counter(), collections.Counter()
    for ln, r in zip(lines, body):
        inst[ln] += int(r[ci])
        samp[ln] += int(r[cs])
    ti, ts = sum(inst.values()), sum(samp.values())
    src = open(os.path.join(ROOT, "snake-slam_b200", "csrc", "orb_kernels.cuh")).read().splitlines()
    print("total warp instructions %d, samples %d" % (ti, ts))
    for ln, n in sorted(inst.items(), key=lambda kv: -kv[1])[:top]:
        text = src[ln[1] - 1].strip()[:90] if ln and ln[0] == "orb_kernels.cuh" else ""
        print("%5.1f%% inst %5.1f%% stall  %s:%d  %s" % (100.0 * n / ti, 100.0 * samp[ln] / max(ts, 1), ln[0], ln[1], text))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)
