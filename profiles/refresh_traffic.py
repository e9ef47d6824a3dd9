"""Rewrite profiles/traffic.json (the DRAM bytes per launch that bench.py reports as roofline.traffic) from an
ncu --set full report.  Usage: python profiles/refresh_traffic.py file.ncu-rep "source text"."""
import csv
import io
import json
import os
import subprocess
import sys

rep, source = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
col = {k: i for i, k in enumerate(hdr)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
want = {"refine_fused": "refine_fused_kernel", "iter_proj_fused": "iter_proj_fused_kernel", "match_prep": "match_prep",
        "gn_accum_stream_kernel<rays>": "gn_accum_stream_kernel<0, 0>"}
out = {"source": source}
for key, pat in want.items():
    best = None
    for r in rows[2:]:
        if pat in r[col["Kernel Name"]]:
            b = sum(float(r[col[m]].replace(",", "")) * scale.get(units[col[m]], 1)
                    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            best = b if best is None else max(best, b)  # the largest launch = the bench-size one
    if best is not None:
        out[key] = best
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json")
json.dump(out, open(path, "w"), indent=1)
print(json.dumps(out, indent=1))
