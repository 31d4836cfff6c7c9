"""python profiles/sass_lines.py <report.ncu-rep> <cubin> <mangled kernel name> [n]
Joins the per-SASS-instruction counters of an ncu report (--import-source on) with nvdisasm's inline line info of the same build
and prints executed warp-instructions / stall samples by (a) call site in the kernel body, (b) line inside the function the
kernel calls directly."""
import csv, subprocess, sys, collections, io, re
def main():
    rep, cubin, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
    dis = subprocess.run(['nvdisasm', '-gi', '-c', cubin], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(dis) if l.startswith('.text.' + kern + ':'))
    locs, chain, fresh = [], [], True
    for l in dis[start + 1:]:
        if l.startswith('\t.section') or l.startswith('//-----'): break
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            if fresh: chain = []; fresh = False
            chain.append((m.group(1).split('/')[-1], int(m.group(2))))
            continue
        m = re.match(r'\s*/\*([0-9a-f]+)\*/\s+(.*?);', l)
        if m:
            fresh = True
            toks = m.group(2).split()
            if toks[0].startswith('@'): toks = toks[1:]
            locs.append((int(m.group(1), 16), toks[0], tuple(chain)))
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
    hdr = rows[hi]
    ci, cs, cn = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
    data = [r for r in rows[hi + 1:] if len(r) > ci and r[0].startswith('0x')]
    base = int(data[0][0], 16)
    bymap = {o: (op, ch) for o, op, ch in locs}
    site, inner, mism = collections.Counter(), collections.Counter(), 0
    ssite, sinner = collections.Counter(), collections.Counter()
    for r in data:
        off = int(r[0], 16) - base
        toks = r[cs].split()
        if toks and toks[0].startswith('@'): toks = toks[1:]
        op, ch = bymap.get(off, ('?', ()))
        if not toks or toks[0].rstrip(';') != op: mism += 1
        v, s = int(r[ci]), int(r[cn] or 0)
        k1 = ch[-1] if ch else ('?', 0)
        k2 = ch[-2] if len(ch) > 1 else k1
        site[k1] += v; ssite[k1] += s; inner[(k1, k2)] += v; sinner[(k1, k2)] += s
    tot, ts = sum(site.values()), sum(ssite.values())
    print(f'{len(data)} SASS instructions, {mism} opcode mismatches; total {tot:.4e} warp-instructions, {ts} samples')
    print('--- by call site in the kernel body')
    for k, v in site.most_common(25): print(f'{100 * v / tot:6.2f}%  samples {100 * ssite[k] / ts:6.2f}%  {k[0]}:{k[1]}')
    print('--- by line one inline level down')
    for k, v in inner.most_common(top):
        print(f'{100 * v / tot:6.2f}%  samples {100 * sinner[k] / ts:6.2f}%  {k[0][0]}:{k[0][1]} <- {k[1][0]}:{k[1][1]}')
main()
