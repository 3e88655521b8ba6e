#!/usr/bin/env python
"""sass_profile.py — attribute an ncu per-instruction profile to source lines, through inlining.

  ncu -i X.ncu-rep --page source --csv --print-source sass > sass.csv
  cuobjdump -xelf all librtx_cuda.so ; nvdisasm -gi -c rtx_render.sm_100a.cubin > gi.txt
  python tools/sass_profile.py sass.csv gi.txt <mangled kernel name> [--by outer|inner|chain|op]

The ncu SASS page lists instructions in address order; nvdisasm -gi lists the same instructions with
`//## File "...", line N inlined at ...` chains.  Joined by ordinal position.
"""
import collections
import csv
import re
import sys


def load_ncu(path):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    i_s, i_e, i_t, i_m = (hdr.index(k) for k in ("Source", "Instructions Executed", "Thread Instructions Executed",
                                                   "# Samples"))
    return [(r[i_s].strip(), int(r[i_e] or 0), int(r[i_t] or 0), int(r[i_m] or 0)) for r in rows[2:] if len(r) > i_t]


def load_gi(path, kernel):
    """-> list of (opcode text, [(file, line) innermost..outermost])"""
    out, chain, inside = [], [], False
    pat = re.compile(r'//## File "([^"]+)", line (\d+)')
    new_chain = True
    for line in open(path):
        if line.startswith(kernel + ":"):
            inside = True
            continue
        if not inside:
            continue
        if line.startswith("\t.section") or line.startswith("//-----"):
            if out:
                break
        m = pat.search(line)
        if m:
            if new_chain:
                chain = []
                new_chain = False
            chain.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m2 = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m2:
            out.append((m2.group(2).strip(), list(chain)))
            new_chain = True
    return out


def main():
    ncu_csv, gi, kernel = sys.argv[1:4]
    by = "outer"
    if "--by" in sys.argv:
        by = sys.argv[sys.argv.index("--by") + 1]
    prof = load_ncu(ncu_csv)
    dis = load_gi(gi, kernel)
    n = min(len(prof), len(dis))
    if len(prof) != len(dis):
        print(f"# warning: {len(prof)} profiled vs {len(dis)} disassembled instructions; joining first {n}",
              file=sys.stderr)
    tot_e = sum(p[1] for p in prof) or 1
    tot_s = sum(p[3] for p in prof) or 1
    agg = collections.defaultdict(lambda: [0, 0, 0])
    for (src, e, t, s), (op, chain) in zip(prof[:n], dis[:n]):
        if not chain:
            key = "?"
        elif by == "outer":
            key = "%s:%d" % chain[-1]
        elif by == "inner":
            key = "%s:%d" % chain[0]
        elif by == "op":
            key = re.sub(r"^@!?U?P\d+\s+", "", src).split()[0]
        else:
            key = " < ".join("%s:%d" % c for c in chain)
        a = agg[key]
        a[0] += e
        a[1] += t
        a[2] += s
    print(f"# total warp instructions {tot_e}, samples {tot_s}")
    for key, (e, t, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:60]:
        if e == 0:
            continue
        print(f"{e / tot_e * 100:6.2f}% inst  {s / tot_s * 100:6.2f}% samples  lanes {t / e:5.1f}  {key}")


if __name__ == "__main__":
    main()
