"""Join an ncu SASS source-page CSV with nvdisasm line info: samples and executed instructions per source line.

    ncu -i rep --page source --csv > sass.csv ; nvdisasm -g -c cubin > dis.txt ; python tools/ncu_lines.py sass.csv dis.txt <kernel substring>
"""
import csv, re, sys, collections
sass_csv, dis, kern = sys.argv[1], sys.argv[2], sys.argv[3]
# --- line info per address of the wanted function
addr_line = {}
cur = None
infun = False
for ln in open(dis):
    if ln.startswith("\t.section") or ".text." in ln and ln.strip().startswith(".section"):
        infun = kern in ln
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)), "inlined" in m.group(3))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', ln)
    if m and infun:
        addr_line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(sass_csv)))
h = rows[1]
ia, isamp, iexec = h.index("Address"), h.index("# Samples"), h.index("Instructions Executed")
stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
base = None
per = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot_s = tot_e = 0
for r in rows[2:]:
    if len(r) < len(h): continue
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if base is None: base = a
    key = addr_line.get(a - base, ("?", 0, False))
    s, e = int(r[isamp] or 0), int(r[iexec] or 0)
    per[key[:2]][0] += s; per[key[:2]][1] += e
    for i in stall_cols:
        v = int(r[i] or 0)
        if v: per[key[:2]][2][h[i][6:]] += v
    tot_s += s; tot_e += e
print(f"total samples {tot_s}, instructions executed {tot_e}")
steps = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
for key, (s, e, st) in sorted(per.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    if s * 200 < tot_s and e * 200 < tot_e: continue
    top = ", ".join(f"{k}:{v}" for k, v in st.most_common(3))
    print(f"{key[0]:18s} {key[1]:5d}  samples {100.0*s/tot_s:5.1f}%  inst/step {e/steps:7.1f}   {top}")
