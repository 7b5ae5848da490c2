"""ncu launch list (--metrics gpu__time_duration.sum --csv --log-file X) -> per-kernel summary CSV on stdout.
python tools/launch_summary.py gpurun_out/launches.csv "header comment" > profiles/rNN_launches_summary.csv"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ik]).replace("tsm::", "")
    tot[name] += float(r[iv].replace(",", "")) * scale[r[iu]]
    cnt[name] += 1
total = sum(tot.values())
if len(sys.argv) > 2:
    print("# " + sys.argv[2])
print("kernel,launches,total_us,avg_us,share")
for k in sorted(tot, key=tot.get, reverse=True):
    print(f'"{k}",{cnt[k]},{tot[k]:.1f},{tot[k] / cnt[k]:.2f},{tot[k] / total:.4f}')
