#!/usr/bin/env python
"""Per-source-line cost of one kernel from an ncu report (development tool).

ncu's CSV source page is SASS-only; nvdisasm -g knows which file:line each SASS instruction
came from (the library is built with -lineinfo). This joins the two by instruction offset.

  python tools/ncu_lines.py gpurun_out/ncu_transport.ncu-rep k_transport_imcILi3ELi4ELb0 [top_n]
"""
import collections
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_lines(section_pat):
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "jaybenne_b200", "libjaybenne_b200.so")],
                          cwd=tmp, stdout=subprocess.DEVNULL)
    cubin = [f for f in glob.glob(os.path.join(tmp, "jb_kernels*.cubin"))][0]
    text = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
    out, inside, loc = {}, False, ("?", 0)
    for line in text:
        if line.startswith("\t.section\t.text."):
            inside = section_pat in line
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "(.*)", line (\d+)', line)
        if m:
            loc = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            out[int(m.group(1), 16)] = (loc, m.group(2).strip())
    return out


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
    by_samples = len(sys.argv) > 4 and sys.argv[4] == "samples"   # rank the lines by stall samples instead
    lines = sass_lines(pat)
    raw = (open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout)
    rows = list(csv.reader(raw.splitlines()))
    hdr = next(r for r in rows if r and r[0] == "Address")
    body = [r for r in rows if len(r) == len(hdr) and r[0].startswith("0x")]
    base = int(body[0][0], 16)
    col = {h: i for i, h in enumerate(hdr)}
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])
    ops = collections.defaultdict(lambda: [0, 0])
    tot = [0, 0, 0, 0]
    for r in body:
        off = int(r[0], 16) - base
        loc, ins = lines.get(off, (("?", 0), r[1]))
        ie, te, smp = int(r[col["Instructions Executed"]]), int(r[col["Thread Instructions Executed"]]), int(r[col["# Samples"]])
        a = agg[loc]
        a[0] += ie; a[1] += te; a[2] += smp; a[3] += 1
        tot[0] += ie; tot[1] += te; tot[2] += smp
        op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
        op = op.split(".")[0]
        ops[op][0] += ie; ops[op][1] += te
    print(f"total warp-inst {tot[0]:.4g}  thread-inst {tot[1]:.4g}  avg active {tot[1] / max(tot[0], 1):.2f}  samples {tot[2]}")
    print("\nby source line (share of warp instructions, of thread instructions, avg active lanes, share of stall samples)")
    src_cache = {}
    for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][2 if by_samples else 0])[:top]:
        f = loc[0]
        if f not in src_cache:
            cand = glob.glob(os.path.join(ROOT, "jaybenne_b200", "csrc", "**", f), recursive=True)
            src_cache[f] = open(cand[0]).read().splitlines() if cand else []
        text = src_cache[f][loc[1] - 1].strip()[:80] if 0 < loc[1] <= len(src_cache[f]) else ""
        print(f"{f[-22:]:>22}:{loc[1]:<4} wi {100 * a[0] / tot[0]:5.2f}%  ti {100 * a[1] / tot[1]:5.2f}%  act {a[1] / max(a[0], 1):4.1f}  "
              f"smp {100 * a[2] / max(tot[2], 1):5.2f}%  n={a[3]:<3} | {text}")
    print("\nby opcode")
    for op, a in sorted(ops.items(), key=lambda kv: -kv[1][0])[:25]:
        print(f"{op:>10} wi {100 * a[0] / tot[0]:5.2f}%  act {a[1] / max(a[0], 1):4.1f}")


if __name__ == "__main__":
    main()
