"""python profiles/ncu_table.py <report.ncu-rep> : one line per kernel of an ncu --set full report (selected raw metrics)"""
import csv, subprocess, sys
COLS = [('gpu__time_duration.sum', 'time'), ('launch__grid_size', 'grid'), ('launch__block_size', 'block'),
        ('launch__registers_per_thread', 'regs'), ('smsp__inst_executed.sum', 'warp-inst'),
        ('smsp__issue_active.avg.pct', 'issue %'), ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps %'),
        ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor %'),
        ('sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'xu %'),
        ('dram__bytes_read.sum', 'dram rd'), ('dram__bytes_write.sum', 'dram wr'),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smem conflicts'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'long sb'),
        ('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'short sb'),
        ('smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'barrier')]


def main():
    out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print('| kernel | ' + ' | '.join(n for _, n in COLS) + ' |')
    print('|---|' + '---|' * len(COLS))
    for r in rows[2:]:
        name = r[idx['Kernel Name']].split('(')[0][:40]
        cells = []
        for m, _ in COLS:
            if m in idx:
                v, u = r[idx[m]], units[idx[m]]
                try:
                    f = float(v.replace(',', ''))
                    v = f'{f:.3g}' if abs(f) < 1e6 else f'{f:.3e}'
                except ValueError:
                    pass
                cells.append(f'{v} {u}'.strip())
            else:
                cells.append('-')
        print(f'| `{name}` | ' + ' | '.join(cells) + ' |')


main()
