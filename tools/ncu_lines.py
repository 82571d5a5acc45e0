#!/usr/bin/env python3
"""Source-line attribution of an `ncu --set full --import-source on` capture without the GUI.

    python tools/ncu_lines.py <report.ncu-rep> <kernel-substring> [--so pathtrace_b200/libb200pt.so] [--top 40]

Joins ncu's per-SASS-instruction counters (source page, SASS view) with the file:line of every instruction taken
from `nvdisasm -g` on the cubins of the SAME build of the library (-lineinfo), by instruction offset, and prints
the executed-instruction and stall-sample shares per source line and per source file.
"""
import argparse
import collections
import csv
import glob
import os
import re
import subprocess
import tempfile


def disasm_lines(so, kernel):
    """[(offset, file, line)] for the first function whose mangled name contains `kernel`."""
    out = []
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, capture_output=True)
        for cubin in sorted(glob.glob(os.path.join(d, "*.cubin"))):
            text = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
            inside, cur = False, ("?", 0)
            for ln in text.splitlines():
                if ln.startswith("\t.section\t.text."):
                    if inside and out:
                        return out
                    inside = kernel in ln
                    continue
                if not inside:
                    continue
                m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
                if m:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                    continue
                m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
                if m:
                    out.append((int(m.group(1), 16), cur[0], cur[1], m.group(2).strip()))
            if out:
                return out
    return out


def ncu_sass(rep, kernel):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    secs, cur = [], None
    for r in csv.reader(raw.splitlines()):
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            secs.append(cur)
        elif cur is not None and r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
            cur["rows"].append(r)
    hits = [s for s in secs if kernel in s["name"]]
    if not hits:
        raise SystemExit("kernel not found in report; have: " + ", ".join(x["name"][:70] for x in secs))
    return hits


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("kernel", help="substring of the MANGLED name (nvdisasm side), e.g. k_neeILb0ELb0")
    ap.add_argument("--ncu", default=None, help="substring of the demangled name in the report, e.g. "
                                                "'k_nee<(bool)0, (bool)0>' (default: derived from the mangled one)")
    ap.add_argument("--so", default="pathtrace_b200/libb200pt.so")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--index", type=int, default=0, help="which captured launch of that kernel")
    a = ap.parse_args()
    dis = disasm_lines(a.so, a.kernel)
    ncu_name = a.ncu
    if ncu_name is None:
        m = re.match(r"([a-z_]+)((?:ILb[01]E?)*(?:Lb[01]E?)*)", a.kernel)
        flags = re.findall(r"Lb([01])", a.kernel)
        ncu_name = m.group(1) + ("<" + ", ".join(f"(bool){f}" for f in flags) + ">" if flags else "")
    sec = ncu_sass(a.rep, ncu_name)[a.index]
    idx = {n: i for i, n in enumerate(sec["hdr"])}
    rows = sec["rows"]
    if len(rows) != len(dis):
        print(f"WARNING: ncu has {len(rows)} instructions, nvdisasm {len(dis)}: library and report differ?")
    base = int(rows[0][idx["Address"]], 16)
    by_off = {off: (f, l, txt) for off, f, l, txt in dis}
    tot_ex = sum(int(r[idx["Instructions Executed"]]) for r in rows)
    tot_sm = sum(int(r[idx["# Samples"]]) for r in rows)
    stall_cols = [n for n in sec["hdr"] if n.startswith("stall_") and "Not Issued" not in n]
    line = collections.defaultdict(lambda: [0, 0, 0, collections.Counter(), 0])
    files = collections.defaultdict(lambda: [0, 0, 0])
    for r in rows:
        off = int(r[idx["Address"]], 16) - base
        f, l, _ = by_off.get(off, ("?", 0, ""))
        ex, sm = int(r[idx["Instructions Executed"]]), int(r[idx["# Samples"]])
        th = int(r[idx["Thread Instructions Executed"]])
        e = line[(f, l)]
        e[0] += ex
        e[1] += sm
        e[2] += th
        e[4] += 1
        for c in stall_cols:
            e[3][c[6:]] += int(r[idx[c]])
        files[f][0] += ex
        files[f][1] += sm
        files[f][2] += th
    print(f"kernel {sec['name'][:90]}\n{len(rows)} SASS instructions, {tot_ex} warp instructions executed, "
          f"{tot_sm} stall samples\n")
    print("per file: exec share, sample share, active lanes per instruction")
    for f, (ex, sm, th) in sorted(files.items(), key=lambda kv: -kv[1][1]):
        print(f"  {f:22s} {ex / tot_ex:6.3f} {sm / max(1, tot_sm):6.3f} {th / max(1, ex):5.1f}")
    print(f"\ntop {a.top} source lines by stall samples: file:line  sass  exec  samples  lanes  main stalls")
    for (f, l), (ex, sm, th, st, n) in sorted(line.items(), key=lambda kv: -kv[1][1])[:a.top]:
        top = ", ".join(f"{k} {v / max(1, sm):.2f}" for k, v in st.most_common(3))
        print(f"  {f + ':' + str(l):28s} {n:4d} {ex / tot_ex:6.3f} {sm / max(1, tot_sm):6.3f} {th / max(1, ex):5.1f}  {top}")


if __name__ == "__main__":
    main()
