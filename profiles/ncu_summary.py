"""Print the handful of ncu metrics the roofline discussion uses from a .ncu-rep file.

    python profiles/ncu_summary.py gpurun_out/<file>.ncu-rep
"""
import csv
import subprocess
import sys

KEEP = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic', 'gpu__time_duration.sum', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'lts__t_bytes.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor.sum', 'smsp__inst_executed.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__cycles_elapsed.avg.per_second',
        'sm__cycles_elapsed.max', 'smsp__cycles_active.avg']


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('---')
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                print(f'{k} [{units[i]}] = {r[i]}')
        for i, h in enumerate(hdr):
            if 'pipe_tensor' in h and h not in KEEP:
                print(f'{h} [{units[i]}] = {r[i]}')


if __name__ == '__main__':
    main(sys.argv[1])
