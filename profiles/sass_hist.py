"""python profiles/sass_hist.py <report.ncu-rep> [n] : executed warp-instructions and stall samples of the first kernel of an ncu
report (--import-source on), grouped by SASS opcode."""
import csv, subprocess, sys, collections, io
def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
    hdr = rows[hi]
    ci, cs, cn = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
    ex, sm = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) <= ci or not r[ci].isdigit(): continue
        toks = r[cs].split()
        if toks and toks[0].startswith('@'): toks = toks[1:]
        op = toks[0].rstrip(';') if toks else '?'
        op = '.'.join(op.split('.')[:2])
        ex[op] += int(r[ci]); sm[op] += int(r[cn] or 0)
    tot, tots = sum(ex.values()), sum(sm.values())
    print(f'total warp-instructions {tot:.4e}, samples {tots}')
    for op, v in ex.most_common(top):
        print(f'{op:28s} {v:14d} {100 * v / tot:6.2f} %   samples {100 * sm[op] / max(tots, 1):6.2f} %')
main()
