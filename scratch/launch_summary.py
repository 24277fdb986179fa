"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import csv, re, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == 'ID'][0]
h, data = rows[hdr], rows[hdr + 1:]
ik, iv, iu = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
agg = collections.OrderedDict()
for r in data:
    name = re.sub(r'\(.*', '', r[ik])[:100]
    v = float(r[iv].replace(',', ''))
    ms = v / 1e6 if r[iu] in ('ns', 'nsecond') else v / 1e3 if r[iu] in ('us', 'usecond') else v
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ms
tot = sum(a[1] for a in agg.values())
print(f"{len(data)} launches, {tot:.3f} ms of kernel time (per-launch times under ncu: cold caches, serialised)")
for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
    print(f"{ms:10.3f} ms {100 * ms / tot:5.1f} %  x{n:4d}  {k}")
