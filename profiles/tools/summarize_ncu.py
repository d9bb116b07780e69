"""Turn an .ncu-rep (ncu --set full) into a small text summary + traffic numbers.

    python profiles/tools/summarize_ncu.py gpurun_out/x.ncu-rep [more.ncu-rep ...] > profiles/x.txt

Runs where there is no GPU (ncu -i only reads the report)."""
import csv
import collections
import json
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__waves_per_multiprocessor',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio']


def to_bytes(v, unit):
    v = float(v.replace(',', ''))
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12}[unit]


def main():
    traffic = {}
    for rep in sys.argv[1:]:
        raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units = rows[0], rows[1]
        for r in rows[2:]:
            name = r[hdr.index('Kernel Name')]
            print('==', rep, '::', name)
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    print('  %-82s %s %s' % (k, r[i], units[i]))
            ir, iw = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
            traffic[name.split('(')[0].strip()] = to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])
        src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(src.splitlines()))
        hdr = rows[1]
        ia, it, isrc = hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('Source')
        byop, thr, tot = collections.Counter(), collections.Counter(), 0
        for r in rows[2:]:
            try:
                n, t = int(r[ia]), int(r[it])
            except ValueError:
                continue
            toks = r[isrc].split()
            op = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
            byop[op] += n
            thr[op] += t
            tot += n
        print('  warp instructions by opcode (share, mean active lanes):')
        for op, n in byop.most_common(14):
            print('    %-8s %5.1f%%  %4.1f' % (op, 100.0 * n / max(tot, 1), thr[op] / max(n, 1)))
    print('TRAFFIC_JSON', json.dumps(traffic))


if __name__ == '__main__':
    main()
