"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.
usage: python profiles/ncu_lines.py src_cs.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
hdr = None; cur_file = None; agg = {}
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if r[0] == 'Function Name' or hdr is None or len(r) < 10: continue
    try: ln = int(r[0])
    except ValueError: continue
    oS = len(hdr) - hdr.index('# Samples'); oI = len(hdr) - hdr.index('Instructions Executed')
    def num(x):
        try: return float(x)
        except ValueError: return 0.0
    s, i = num(r[len(r) - oS]), num(r[len(r) - oI])
    a = agg.setdefault((cur_file, ln, r[1][:100]), [0, 0]); a[0] += s; a[1] += i
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print('samples', ts, 'warp-instructions', ti)
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]:>16}:{k[1]:<4} stall={a[0]/ts*100:5.1f}% inst={a[1]/ti*100:5.1f}%  {k[2]}")
