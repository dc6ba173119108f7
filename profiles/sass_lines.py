"""Joins an `ncu --page source --print-source sass --csv` export with `nvdisasm -g -c` output of the same cubin to
attribute executed instructions to source lines.  usage: sass_lines.py sass.csv dis.txt <kernel substring> [top]"""
import collections
import csv
import re
import sys

sass_csv, dis_txt, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25

# --- disassembly: per kernel, list of (file:line, text)
dis = collections.OrderedDict()
cur = None
loc = "?"
for ln in open(dis_txt, errors="replace"):
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        cur = dis.setdefault(m.group(1), [])
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        loc = f"{m.group(1).split('/')[-1]}:{m.group(2)}"
        continue
    if cur is not None and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        cur.append((loc, ln.strip()))

rows = list(csv.reader(open(sass_csv)))
kern = []
k = None
for r in rows:
    if r and r[0] == "Kernel Name":
        k = {"name": r[1], "rows": []}
        kern.append(k)
    elif r and r[0] == "Address":
        k["hdr"] = r
    elif k is not None and r:
        k["rows"].append(r)
seen = set()
for k in kern:
    if kname not in k["name"] or k["name"] in seen:
        continue
    seen.add(k["name"])
    ix = {n: i for i, n in enumerate(k["hdr"])}
    # find the disassembly with the same instruction count
    cands = [(n, v) for n, v in dis.items() if len(v) == len(k["rows"])]
    if not cands:
        print("no disassembly of matching length for", k["name"], len(k["rows"]), [(n[:40], len(v)) for n, v in dis.items()])
        continue
    name, ins = cands[0]
    agg = collections.defaultdict(lambda: [0, 0, 0])
    tot = 0
    for (loc, text), r in zip(ins, k["rows"]):
        wi = int(r[ix["Instructions Executed"]])
        ti = int(r[ix["Thread Instructions Executed"]])
        sm = int(r[ix["# Samples"]])
        a = agg[loc]
        a[0] += wi
        a[1] += ti
        a[2] += sm
        tot += wi
    tots = sum(a[2] for a in agg.values())
    print(f"{k['name'][:70]}: {tot} warp instructions, {len(ins)} SASS instructions")
    for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"  {loc:28s} inst {a[0] / tot * 100:5.1f}%  lanes {a[1] / max(a[0], 1):5.1f}  stall-samples {a[2] / max(tots, 1) * 100:5.1f}%")
