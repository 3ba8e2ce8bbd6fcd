"""Hot instructions of an `ncu --page source --csv --print-source sass` export: opcode histogram of the instructions executed at
>= frac of the maximum count, and the top stall-sample lines.  python scripts/ncu_hot.py file_sass.csv [frac] [ntop]"""
import csv, sys, collections, re
f = sys.argv[1]; frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5; ntop = int(sys.argv[3]) if len(sys.argv) > 3 else 15
rows = list(csv.reader(open(f)))
hdr = rows[1]
iS, iE, iN = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
iW, iWi = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Ideal")
body = [r for r in rows[2:] if len(r) > iW and r[iE].isdigit()]
mx = max(int(r[iE]) for r in body)
tot = sum(int(r[iE]) for r in body)
hot = [r for r in body if int(r[iE]) >= frac * mx]
c = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", r[iS].strip()).split()[0].split(".")[0] for r in hot)
print("max count %d, total executed %d, hot instructions %d (%.1f%% of executed)" % (mx, tot, len(hot), 100.0 * sum(int(r[iE]) for r in hot) / tot))
print(dict(c))
print("shared wavefronts: %d (ideal %d)" % (sum(int(r[iW]) for r in body), sum(int(r[iWi]) for r in body)))
ts = sum(int(r[iN]) for r in body)
for r in sorted(body, key=lambda r: -int(r[iN]))[:ntop]:
    st = [(int(r[i]), hdr[i]) for i in range(len(hdr)) if hdr[i].startswith("stall_") and "Not Issued" not in hdr[i] and r[i].isdigit() and int(r[i]) > 0]
    st.sort(reverse=True)
    print("%5.2f%% x%-10s %-60s %s" % (100.0 * int(r[iN]) / ts, r[iE], r[iS].strip()[:60], " ".join("%s:%d" % (h[6:], v) for v, h in st[:3])))
