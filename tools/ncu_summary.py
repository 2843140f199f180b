"""Compact text summary of an .ncu-rep (raw page): duration, occupancy, issue/pipe utilisation, memory traffic and
the warp-stall breakdown per kernel launch.   python tools/ncu_summary.py report.ncu-rep > profiles/xxx.txt"""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_static', 'launch__waves_per_multiprocessor', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__throughput.avg.pct_of_peak_sustained_elapsed']


def main():
    rep = sys.argv[1]
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    print('# %s' % rep)
    for r in data:
        print('\n## launch %s  %s' % (r[ix['ID']], r[ix['Kernel Name']][:110]))
        for k in KEYS:
            if k in ix:
                print('%-72s %s %s' % (k, r[ix[k]], units[ix[k]]))
        st = []
        for h in hdr:
            if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio'):
                try:
                    st.append((float(r[ix[h]]), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]))
                except ValueError:
                    pass
        print('warp stalls per issued instruction: ' + ', '.join('%s %.2f' % (n, v) for v, n in sorted(st, reverse=True)[:8]))


if __name__ == '__main__':
    main()
