"""Condenses an ncu report (ncu --set full ...) into one CSV row per distinct kernel (first launch after warm-up).

    python profiles/summarize_ncu.py gpurun_out/prof_r1e.ncu-rep profiles/r01_ncu_summary.csv [skip_launches_per_kernel]
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_read.sum",
    "lts__t_sectors_op_write.sum",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    skip = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    seen = {}
    picked = []
    for r in rows[2:]:
        name = r[ki]
        short = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        n = seen.get(short, 0)
        seen[short] = n + 1
        if n == skip or (n == 0 and short.startswith(("occ_grad", "adam", "count_opaque", "build_rays"))):
            picked.append((short, r))
    cols = [m for m in METRICS if m in hdr]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{m} [{units[hdr.index(m)]}]" for m in cols] + ["dram_GBps", "launches_in_capture"])
        for short, r in picked:
            vals = [r[hdr.index(m)] for m in cols]
            try:
                t = float(r[hdr.index("gpu__time_duration.sum")].replace(",", ""))
                tu = units[hdr.index("gpu__time_duration.sum")]
                t_s = t * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(tu, 1e-9)
                b = 0.0
                for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    v = float(r[hdr.index(m)].replace(",", ""))
                    u = units[hdr.index(m)]
                    b += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                gbps = f"{b / t_s / 1e9:.0f}"
            except Exception:
                gbps = ""
            w.writerow([short] + vals + [gbps, seen[short]])
    print(open(out).read())


if __name__ == "__main__":
    main()
