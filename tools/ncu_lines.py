#!/usr/bin/env python
"""Per-source-line summary of an ncu capture: joins `ncu --page source --csv` (SASS rows) with
`nvdisasm -g` line info of the same cubin, in instruction order.

    [NCU_LAUNCH=k] python tools/ncu_lines.py REPORT.ncu-rep KERNEL_SUBSTRING [top_n]

Prints, per source line: warp instructions executed, average active threads, stall samples.
"""
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_lines(kernel_sub):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.environ.get("CRT_LIB", os.path.join(ROOT, "computeraytracer_b200", "libcrt.so"))], cwd=tmp,
                   check=True, capture_output=True)
    out = ""
    for f in sorted(os.listdir(tmp)):
        if f.endswith(".cubin"):
            out += subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    cur, infn, res = None, False, []
    for ln in out.splitlines():
        if ln.startswith(".text."):
            infn = kernel_sub in ln
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            res.append((int(m.group(1), 16), m.group(2).strip(), cur))
    return res


def main():
    rep, ksub = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    # find the block of the wanted kernel
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            blocks.append(cur)
        elif cur is not None:
            cur["rows"].append(r)
    which = int(os.environ.get("NCU_LAUNCH", "0"))   # n-th captured launch of the kernel
    blk = [b for b in blocks if ksub in b["name"].replace("(bool)0", "Lb0").replace("(bool)1", "Lb1") or ksub in b["name"]][which]
    hdr = blk["rows"][0]
    data = blk["rows"][1:]
    ci = {n: i for i, n in enumerate(hdr)}
    m_ = re.match(r"(\w+)<(.*)>", ksub)
    mangled = (m_.group(1) + "I" + "".join("Lb" + a.strip().replace("(bool)", "") + "E" for a in m_.group(2).split(","))) if m_ else ksub
    sass = sass_lines(mangled if "Lb" in mangled else "pt_megakernelILb0" if "megakernel" in ksub else mangled)
    if len(sass) != len(data):
        print(f"warning: {len(sass)} disassembled instructions vs {len(data)} profiled rows", file=sys.stderr)
    agg = {}
    tot_i = tot_t = tot_s = 0
    for k, r in enumerate(data):
        if k >= len(sass):
            break
        key = sass[k][2]
        ie = int(r[ci["Instructions Executed"]] or 0)
        te = int(r[ci["Thread Instructions Executed"]] or 0)
        ss = int(r[ci["# Samples"]] or 0)
        a = agg.setdefault(key, [0, 0, 0])
        a[0] += ie; a[1] += te; a[2] += ss
        tot_i += ie; tot_t += te; tot_s += ss
    print(f"total warp inst {tot_i:,}  avg threads {tot_t / max(tot_i, 1):.2f}  samples {tot_s:,}")
    src_cache = {}
    for key, (ie, te, ss) in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
        text = ""
        if key:
            p = os.path.join(ROOT, "computeraytracer_b200", "csrc", key[0])
            if not os.path.exists(p):
                p = os.path.join(ROOT, "include", key[0])
            if os.path.exists(p):
                src_cache.setdefault(p, open(p).read().splitlines())
                if key[1] - 1 < len(src_cache[p]):
                    text = src_cache[p][key[1] - 1].strip()[:90]
        print(f"{100 * ss / max(tot_s, 1):5.1f}% smp  {100 * ie / max(tot_i, 1):5.1f}% inst  thr {te / max(ie, 1):5.1f}  {key[0] if key else '?'}:{key[1] if key else 0:<4d} {text}")


if __name__ == "__main__":
    main()
