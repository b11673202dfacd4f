"""Summarise an .ncu-rep (read here with `ncu -i`) into a small CSV kept under profiles/.

usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/r1_x_summary.csv
One row per captured launch; only the metrics DESIGN.md / bench.py quote."""
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic",
]


def to_bytes(v, unit):
    mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit)
    return float(v) * mul if mul else None


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{m} [{units[col[m]]}]" for m in METRICS if m in col] + ["dram_bytes_total"])
        for r in body:
            name = r[col["Kernel Name"]].split("(")[0].replace("void ", "")
            vals = [r[col[m]] for m in METRICS if m in col]
            tot = sum(to_bytes(r[col[m]], units[col[m]]) or 0.0
                      for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            w.writerow([name] + vals + [int(tot)])
    print(open(out).read())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
