"""Summarise an `ncu --page source --csv --print-source sass` export: runs of SASS with equal execution counts
(= loop bodies), their share of issued instructions, lanes per instruction and stall samples.
usage: python tools/hot_sass.py file.csv [min_share_percent]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hdr = rows[1]
data = []
for r in rows[2:]:  # one block per kernel instance in the report: keep the first
    if len(r) != len(hdr): break
    if r[:2] == hdr[:2]: break
    data.append(r)
iS = hdr.index('Source'); iI = hdr.index('Instructions Executed'); iT = hdr.index('Thread Instructions Executed'); iSm = hdr.index('# Samples')
tot = sum(int(r[iI]) for r in data); tots = sum(int(r[iSm]) for r in data)
print('total warp-inst', tot, 'samples', tots, 'sass lines', len(data))
seg = []; cur = None
for k, r in enumerate(data):
    c = int(r[iI])
    if cur and abs(c - cur[2]) <= 0.02 * max(c, cur[2], 1):
        cur[1] = k; cur[3] += c; cur[4] += int(r[iT]); cur[5] += int(r[iSm])
    else:
        if cur: seg.append(cur)
        cur = [k, k, c, c, int(r[iT]), int(r[iSm])]
seg.append(cur)
for s in seg:
    if s[3] > thr / 100 * tot:
        print(f"{s[0]:5d}-{s[1]:5d} n={s[1]-s[0]+1:4d} per={s[2]:>11d} share={s[3]/tot*100:5.1f}% lanes={s[4]/max(s[3],1):5.1f} samples={s[5]/tots*100:5.1f}%")
if len(sys.argv) > 3:
    a, b = map(int, sys.argv[3].split('-'))
    for k in range(a, b + 1):
        r = data[k]
        print(f"{k:5d} {int(r[iI]):>10d} {int(r[iT])/max(int(r[iI]),1):5.1f} {int(r[iSm]):5d}  {r[iS]}")
