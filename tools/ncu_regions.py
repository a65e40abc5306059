"""Warp instructions and stall samples per REGION of a kernel's source (tuning aid).

    ncu -i rep --page source --csv > sass.csv ; nvdisasm -gi -c cubin > dis.txt
    python tools/ncu_regions.py sass.csv dis.txt <kernel substring> <steps> <main file> '[["name", first_line], ...]'

Every SASS instruction is attributed to the line of <main file> at the outer end of its inlining chain."""
import csv, re, sys, collections, json
sass_csv, dis, kern, steps, main = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4]), sys.argv[5]
bounds = json.loads(sys.argv[6])
addr_line = {}; chain = []; infun = False; prev_was_note = False
for ln in open(dis):
    if ln.strip().startswith(".section"): infun = kern in ln
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
    if m:
        if not prev_was_note: chain = []
        chain.append(m.groups()); prev_was_note = True
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', ln)
    if m:
        if infun and chain:
            f, l, cf, cl = chain[-1]
            loc = (cf.split("/")[-1], int(cl)) if cf else (f.split("/")[-1], int(l))
            addr_line[int(m.group(1), 16)] = loc
        prev_was_note = False
rows = list(csv.reader(open(sass_csv))); h = rows[1]
ia, isamp, iexec = h.index("Address"), h.index("# Samples"), h.index("Instructions Executed")
def region(line):
    name = bounds[0][0]
    for n, first in bounds:
        if line >= first: name = n
    return name
tot = collections.defaultdict(lambda: [0, 0]); base = None
for r in rows[2:]:
    if len(r) < len(h): continue
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if base is None: base = a
    key = addr_line.get(a - base)
    name = region(key[1]) if key and key[0] == main else f"other:{key[0] if key else '?'}"
    tot[name][0] += int(r[isamp] or 0); tot[name][1] += int(r[iexec] or 0)
ts = sum(v[0] for v in tot.values()); te = sum(v[1] for v in tot.values())
for n in [b[0] for b in bounds] + sorted(k for k in tot if k.startswith("other:")):
    s, e = tot.get(n, (0, 0))
    print(f"{n:34s} inst/step {e/steps:8.1f}  samples {100*s/ts:5.1f}%")
print("total inst/step", te/steps)
