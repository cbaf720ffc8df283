"""Top SASS / source lines of one profiled kernel by warp-stall samples, from an ncu report captured with
--set full --import-source on (the report itself is too large to keep).

    python profiles/top_source_lines.py report.ncu-rep out.csv [n_lines]
"""
import csv
import io
import subprocess
import sys


def main():
    rep, out = sys.argv[1], sys.argv[2]
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if any("Sampling" in c for c in r))
    hdr = rows[hdr_i]
    col = next(i for i, c in enumerate(hdr) if "Warp Stall Sampling (All" in c)
    keep = [i for i, c in enumerate(hdr) if c in ("Address", "Source") or "Sampling" in c or c.startswith("# ") or "Instructions Executed" in c]
    body = []
    for r in rows[hdr_i + 1:]:
        try:
            body.append((float(r[col].replace(",", "")), r))
        except (ValueError, IndexError):
            pass
    total = sum(v for v, _ in body) or 1.0
    body.sort(key=lambda t: -t[0])
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["share_of_stall_samples_%"] + [hdr[i] for i in keep])
        for v, r in body[:n]:
            w.writerow([f"{100 * v / total:.2f}"] + [r[i] if i < len(r) else "" for i in keep])
    print(open(out).read()[:6000])


if __name__ == "__main__":
    main()
