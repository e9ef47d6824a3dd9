"""Digest an .ncu-rep (via `ncu -i … --page raw --csv`) into one row per kernel with the metrics the
roofline discussion needs.  Usage: python profiles/ncu_digest.py file.ncu-rep [out.md]"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time_us", 1e-3),
    ("dram__bytes_read.sum", "dram_rd_MB", 1e-6),
    ("dram__bytes_write.sum", "dram_wr_MB", 1e-6),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%", 1),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_%", 1),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1_%", 1),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%", 1),
    ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "fmaheavy_%", 1),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu_%", 1),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "fp64_%", 1),
    ("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "dmma_%", 1),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_%", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_%", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
]


def main():
    rep = sys.argv[1]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    col = {k: i for i, k in enumerate(hdr)}
    out = ["| # | kernel | " + " | ".join(k for _, k, _ in KEYS) + " |", "|---|---|" + "---:|" * len(KEYS)]
    for n, r in enumerate(rows[2:]):
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("m3r::", "")[:40]
        vals = []
        for key, _, sc in KEYS:
            if key not in col or r[col[key]] in ("", "n/a"):
                vals.append("-")
                continue
            v = float(r[col[key]].replace(",", ""))
            u = units[col[key]]
            if key.startswith("dram__bytes"):
                v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            if key.startswith("gpu__time"):
                v *= {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(u, 1)
            v *= sc
            vals.append(f"{v:.1f}" if abs(v) < 1e4 and v != int(v) else f"{v:.0f}")
        out.append(f"| {n} | `{name}` | " + " | ".join(vals) + " |")
    s = "\n".join(out) + "\n"
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(s)
    print(s)


main()
