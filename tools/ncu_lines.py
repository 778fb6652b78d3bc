#!/usr/bin/env python
"""Aggregate an ncu report's per-instruction warp-stall samples by CUDA source line.

    python tools/ncu_lines.py <report.ncu-rep> <kernel regex> [--lib mazetd_b200/libmazetd_b200.so] [--top 40]

ncu's CSV source page is SASS-only; line numbers come from `nvdisasm -g` on the cubin inside the .so (compiled with
-lineinfo). Inlined callees are attributed to the innermost line. Prints: samples, share, executed warp
instructions, the dominant stall reason, file:line and the source text.
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(lib, kernel_re):
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
    out = {}
    for f in os.listdir(tmp):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur_fn, cur_line = None, None
        for ln in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                cur_fn = m.group(1)
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur_line = (m.group(1), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m and cur_fn and re.search(kernel_re, cur_fn):
                out.setdefault(cur_fn, {})[int(m.group(1), 16)] = (cur_line, m.group(2).strip())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("kernel")
    ap.add_argument("--lib", default="mazetd_b200/libmazetd_b200.so")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--launch", type=int, default=0, help="which matching launch in the report")
    ap.add_argument("--fn", default=None, help="regex on the MANGLED function name when the kernel is a template with several instantiations (e.g. ILi2E)")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "--kernel-name", "regex:" + a.kernel], capture_output=True, text=True).stdout
    # the CSV holds one block per launch, each starting with a "Kernel Name" row
    blocks = re.split(r'(?m)^"Kernel Name",', raw)[1:]
    if not blocks:
        sys.exit("no kernel matched")
    blk = blocks[min(a.launch, len(blocks) - 1)]
    rows = list(csv.reader(io.StringIO(blk.split("\n", 1)[1])))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    lines = sass_lines(a.lib, a.kernel)
    cands = [k for k in lines if a.fn is None or re.search(a.fn, k)]
    fn = max(cands, key=lambda k: len(lines[k]))
    amap = lines[fn]
    base = None
    agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    total = 0
    for r in rows[1:]:
        if len(r) < len(hdr):
            continue
        addr = int(r[col["Address"]], 16) if r[col["Address"]].startswith("0x") else int(r[col["Address"]])
        if base is None:
            base = addr
        off = addr - base
        s = int(r[col["# Samples"]] or 0)
        ex = int(r[col["Instructions Executed"]] or 0)
        line = amap.get(off, (None, ""))[0]
        e = agg[line]
        e[0] += s
        e[1] += ex
        for h in stall_cols:
            v = int(r[col[h]] or 0)
            if v:
                e[2][h] += v
        total += s
    srcs = {}
    print(f"kernel {fn}: {total} samples")
    for line, (s, ex, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[: a.top]:
        text = ""
        if line:
            if line[0] not in srcs:
                try:
                    srcs[line[0]] = open(line[0]).read().splitlines()
                except Exception:
                    srcs[line[0]] = []
            L = srcs[line[0]]
            text = L[line[1] - 1].strip() if 0 < line[1] <= len(L) else ""
        top = ",".join(f"{k[6:]}:{v}" for k, v in st.most_common(2))
        where = f"{os.path.basename(line[0])}:{line[1]}" if line else "?"
        print(f"{s:8d} {100.0 * s / max(total, 1):5.1f}%  exec={ex:10d}  {top:32s} {where:22s} {text[:100]}")


if __name__ == "__main__":
    main()
