"""Print selected raw metrics of every kernel in an ncu report as a markdown table.

    python profiles/ncu_table.py gpurun_out/x.ncu-rep
"""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_atom.sum', 'lts__t_sectors_op_read.sum',
        'lts__t_sectors_op_write.sum', 'sm__inst_executed_pipe_tensor.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active',
        'l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    cols = [(k, hdr.index(k)) for k in KEYS if k in hdr]
    print("| metric | unit | " + " | ".join(r[name_i].split("(")[0][-40:] for r in data) + " |")
    print("|---|---|" + "---|" * len(data))
    for k, i in cols:
        print(f"| {k} | {units[i]} | " + " | ".join(r[i] for r in data) + " |")


if __name__ == "__main__":
    main()
