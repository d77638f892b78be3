"""Extract the metrics we report from an .ncu-rep into a markdown table.
usage: python tools/ncu_extract.py report.ncu-rep > table.md   (needs ncu on PATH; no GPU needed)"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_fp8_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_uniform.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__cluster_size", "launch__shared_mem_per_block_dynamic",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max.per_second", "sm__cycles_elapsed.max",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        print(f"kernel `{name[:120]}`\n\n| metric | unit | value |\n|---|---|---|")
        for m in WANT:
            if m in hdr:
                i = hdr.index(m)
                print(f"| {m} | {units[i]} | {r[i]} |")
        tens = [h for h in hdr if "tensor" in h and h not in WANT]
        for h in tens[:12]:
            i = hdr.index(h)
            print(f"| {h} | {units[i]} | {r[i]} |")
        print()


if __name__ == "__main__":
    main()
