"""Pipe utilisation, issue rate, stall samples and DRAM bytes of every launch in a report's raw page:
    ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_pipes.py raw.csv   (run here, no GPU)"""
import csv
import re
import sys

PAT = re.compile(r'^(Kernel Name|gpu__time_duration\.sum|sm__inst_executed_pipe_\w+\.avg\.pct_of_peak_sustained_active|'
                 r'sm__pipe_\w+_cycles_active\.avg\.pct_of_peak_sustained_active|smsp__issue_active\.avg\.pct_of_peak_sustained_active|'
                 r'smsp__inst_executed\.sum|sm__warps_active\.avg\.pct_of_peak_sustained_active|sm__cycles_elapsed\.avg|'
                 r'l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum(\.pct_of_peak_sustained_elapsed)?|dram__bytes_read\.sum|dram__bytes_write\.sum|'
                 r'launch__registers_per_thread|smsp__pcsamp_warps_issue_stalled_\w+)$')


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = [i for i, h in enumerate(hdr) if PAT.match(h)]
    for r in rows[2:]:
        print('----')
        stalls = []
        for i in idx:
            v = r[i]
            try:
                f = float(v.replace(',', ''))
            except ValueError:
                f = None
            if f == 0:
                continue
            if hdr[i].startswith('smsp__pcsamp'):
                if 'not_issued' not in hdr[i]:
                    stalls.append((f, hdr[i][len('smsp__pcsamp_warps_issue_stalled_'):]))
                continue
            print(hdr[i], '=', v, units[i])
        tot = sum(f for f, _ in stalls) or 1.0
        print('stall samples:', ', '.join('%s %.1f%%' % (n, 100 * f / tot) for f, n in sorted(stalls, reverse=True)[:8]))


if __name__ == '__main__':
    main(sys.argv[1])
