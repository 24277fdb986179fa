"""Basic-block profile of one kernel from `ncu --page source --csv --print-source sass`: share of warp instructions and lanes per instruction per block."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; iS = hdr.index("Source"); iE = hdr.index("Instructions Executed"); iT = hdr.index("Thread Instructions Executed")
data = []
for r in rows[2:]:
    if len(r) > iT and r[iE].isdigit():
        data.append((r[iS].strip(), int(r[iE]), int(r[iT])))
tot = sum(d[1] for d in data)
print("instr", len(data), "warp inst", tot, "lanes", round(sum(d[2] for d in data) / tot, 2))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
blocks = []; start = 0
for i in range(1, len(data) + 1):
    if i == len(data) or data[i][1] != data[start][1]:
        blocks.append((start, i)); start = i
KEEP = ('LDG', 'STG', 'LDS', 'STS', 'LDL', 'STL', 'MUFU', 'I2F', 'CALL', 'VOTE', 'ATOM', 'RED', 'SHFL', 'BSSY', 'BSYNC', 'BRA', 'WARPSYNC', 'RET')
for a, b in blocks:
    e = sum(d[1] for d in data[a:b]); t = sum(d[2] for d in data[a:b])
    if e / tot < thr: continue
    ops = [(d[0].split()[1] if d[0].startswith('@') else d[0].split()[0]) for d in data[a:b]]
    marks = [o.split('.')[0] for o in ops if o.startswith(KEEP)]
    print(f"{a:5d}-{b:5d} n={b - a:4d} share={e / tot * 100:5.1f}% lanes={t / e if e else 0:5.1f}  {' '.join(marks)[:140]}")
