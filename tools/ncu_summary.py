"""Compact per-kernel table from an ncu report:  python tools/ncu_summary.py <file.ncu-rep> [out.txt]"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "us"), ("sm__cycles_elapsed.max", "cyc"),
        ("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active", "tc_act%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("lts__t_sector_hit_rate.pct", "l2hit%"), ("launch__registers_per_thread", "regs"),
        ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "tc_smem%")]


def main() -> None:
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = {h: i for i, h in enumerate(rows[0])}
    lines = [f"{'kernel':42s} {'grid':>6s} {'block':>6s} " + " ".join(f"{n:>9s}" for _, n in COLS)]
    for r in rows[2:]:
        name = r[hdr["Kernel Name"]].replace("<unnamed>::", "").replace("void ", "")[:42]
        grid = r[hdr["Grid Size"]].strip("()").split(",")[0]
        blk = r[hdr["Block Size"]].strip("()").split(",")[0]
        vals = []
        for m, _ in COLS:
            v = r[hdr[m]] if m in hdr else ""
            try:
                vals.append(f"{float(v.replace(',', '')):9.2f}")
            except ValueError:
                vals.append(f"{v[:9]:>9s}")
        lines.append(f"{name:42s} {grid:>6s} {blk:>6s} " + " ".join(vals))
    out = "\n".join(lines) + "\n"
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(out)
    print(out)


if __name__ == "__main__":
    main()
