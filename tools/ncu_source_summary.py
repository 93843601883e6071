"""Summarise the source page of an ncu report (ncu -i X.ncu-rep --page source --csv > X.csv): stall-reason totals,
samples per address block, and the top instructions.  Usage: python tools/ncu_source_summary.py X.csv [block_hex] [top_n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
base = int(data[0][ix["Address"]], 16)
G = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0x200
TOP = int(sys.argv[3]) if len(sys.argv) > 3 else 25
cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {h: 0 for h in cols}
samples = 0
for r in data:
    samples += int(r[ix["# Samples"]])
    for h in cols:
        tot[h] += int(r[ix[h]])
print("total samples", samples)
for h, v in sorted(tot.items(), key=lambda x: -x[1])[:9]:
    print(f"  {h:26s} {v:8d} {100 * v / max(samples, 1):5.1f}%")
blk = {}
for r in data:
    a = int(r[ix["Address"]], 16) - base
    e = blk.setdefault(a // G, [0, 0])
    e[0] += int(r[ix["# Samples"]])
    e[1] += int(r[ix["Instructions Executed"]])
print("blocks with >= 1% of the samples:")
for b in sorted(blk):
    if blk[b][0] >= samples / 100:
        print(f"  {b * G:#08x} samples {blk[b][0]:6d} executed {blk[b][1]:10d}")
print("top instructions:")
for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:TOP]:
    st = [(h[6:], int(r[ix[h]])) for h in cols if int(r[ix[h]]) > 0]
    st.sort(key=lambda x: -x[1])
    a = int(r[ix["Address"]], 16) - base
    print(f"  {a:#08x} {int(r[ix['# Samples']]):6d} {int(r[ix['Instructions Executed']]):9d}  {r[ix['Source']].strip()[:60]:60s} {st[:3]}")
