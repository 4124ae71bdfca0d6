"""Summarise an ncu report: key raw metrics + the source lines with most stall samples.

    python profiles/ncu_lines.py gpurun_out/x.ncu-rep [top_n] [kernel_index]
"""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_registers']


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    print("Kernel:", [r[hdr.index("Kernel Name")] for r in rows[2:]])
    for i, h in enumerate(hdr):
        if h in KEYS:
            print(f"{h} [{units[i]}]:", [r[i] for r in rows[2:]])
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "source", "--csv",
                                            "--print-source", "cuda,sass"]))))
    his = [i for i, r in enumerate(rows) if r and r[0] == "Line No"]
    if not his:
        return
    hi = his[min(which, len(his) - 1)]
    hdr = rows[hi]
    end = ([h for h in his if h > hi] + [len(rows)])[0]
    samp, inst = hdr.index("# Samples"), hdr.index("Instructions Executed")
    lines = [r for r in rows[hi + 1:end] if r and r[0].isdigit()]
    tot = sum(int(r[samp]) for r in lines if r[samp].isdigit())
    print(f"total samples (kernel #{which}):", tot)
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    lines.sort(key=lambda r: -int(r[samp]) if r[samp].isdigit() else 0)
    for r in lines[:top]:
        st = sorted(((int(r[i]) if r[i].isdigit() else 0, hdr[i][6:]) for i in stall_cols),
                    reverse=True)[:3]
        print(f"{r[0]:>4} {int(r[samp]):5d} {100*int(r[samp])/max(tot,1):5.1f}% inst={r[inst]:>8} "
              f"{r[1].strip()[:64]:<64} {st}")


if __name__ == "__main__":
    main()
