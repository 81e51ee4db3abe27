"""Prints the key metrics of an .ncu-rep (one line per metric, one column per profiled launch).
usage: python profiles/ncu_keys.py report.ncu-rep [extra-substring ...]"""
import csv
import subprocess
import sys

WANT = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_fp64', 'sm__pipe_fp64_cycles_active',
        'smsp__inst_executed.sum', 'sass__inst_executed_local', 'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_write.sum',
        'lts__t_sectors_op_read.sum', 'sm__throughput.avg.pct', 'l1tex__throughput.avg.pct', 'lts__throughput.avg.pct',
        'smsp__issue_active.avg.pct', 'launch__shared_mem_per_block', 'smsp__average_warps_issue_stalled',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared', 'smsp__inst_executed_pipe_lsu', 'launch__waves_per_multiprocessor',
        'sm__cycles_active.avg', 'smsp__cycles_active.avg', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum']


def main():
    rep = sys.argv[1]
    want = WANT + sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    for i, h in enumerate(hdr):
        if any(w in h for w in want):
            vals = [r[i] for r in data]
            if h.startswith('smsp__average_warps_issue_stalled') and all(v in ('0', '0.000000', '') or float(v or 0) < 0.05 for v in vals):
                continue
            print("%-100s %-14s %s" % (h, units[i], "  ".join(vals)))


if __name__ == "__main__":
    main()
