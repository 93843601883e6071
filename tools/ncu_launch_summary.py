"""Per-kernel means from an ncu launch-list CSV (gpu__time_duration.sum [+ dram bytes]), last N steps only.
Usage: python tools/ncu_launch_summary.py FILE.csv [skip_first_fraction]"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if r]
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
acc = defaultdict(lambda: defaultdict(list))
for r in rows[1:]:
    if len(r) < len(hdr):
        continue
    name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").replace("ctcb200::", "")
    if "ctc_" not in name and "fc_" not in name:
        continue
    v = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    if r[ix["Metric Name"]].startswith("gpu__time"):
        v = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    else:
        v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    acc[name][r[ix["Metric Name"]]].append(v)
for name, m in acc.items():
    out = [f"{name:46s} launches {len(next(iter(m.values()))):4d}"]
    for k, vs in m.items():
        vs = vs[len(vs) // 2:]  # warm half
        mean = sum(vs) / len(vs)
        out.append(f"{k.split('__')[-1]}: {mean / 1e6:9.2f} MB" if "bytes" in k else f"time {mean:8.1f} us")
    print("  ".join(out))
