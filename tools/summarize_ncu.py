"""Turn an .ncu-rep (from `ncu --set full`) into the small per-launch table kept under profiles/.
usage: python tools/summarize_ncu.py gpurun_out/prof_r1.ncu-rep > profiles/r1_ncu_full_summary.md"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "launch__occupancy_limit_registers"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    units = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [c for c in WANT if c in idx]
    print("| # | kernel | grid | block | regs | time | DRAM read | DRAM write | DRAM GB/s | DRAM % peak | long-scoreboard stall % |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    for n, r in enumerate(rows[2:]):
        def g(name):
            return r[idx[name]] if name in idx else ""

        def num(name):
            try:
                return float(g(name).replace(",", ""))
            except ValueError:
                return float("nan")
        name = g("Kernel Name")[:70]
        t, tu = num("gpu__time_duration.sum"), units[idx["gpu__time_duration.sum"]]
        t_us = t / 1e3 if tu == "ns" else (t if tu in ("us", "usecond") else t * 1e3 if tu == "ms" else t)
        rd, ru = num("dram__bytes_read.sum"), units[idx["dram__bytes_read.sum"]]
        wr, wu = num("dram__bytes_write.sum"), units[idx["dram__bytes_write.sum"]]
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd_b, wr_b = rd * mult.get(ru, 1), wr * mult.get(wu, 1)
        gbps = (rd_b + wr_b) / (t_us * 1e-6) / 1e9 if t_us == t_us and t_us > 0 else float("nan")
        print(f"| {n} | `{name}` | {g('launch__grid_size')} | {g('launch__block_size')} | {g('launch__registers_per_thread')} | "
              f"{t_us:.1f} us | {rd_b / 1e6:.2f} MB | {wr_b / 1e6:.2f} MB | {gbps:.0f} | "
              f"{g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')} | "
              f"{g('smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct')} |")


if __name__ == "__main__":
    main(sys.argv[1])
