"""Print a per-launch table from an `ncu --csv --metrics ...` log: launches after the last one whose
kernel name contains `anchor` (argv[2]). Diagnostic."""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
anchor = sys.argv[2]
for i, r in enumerate(rows):
    if 'Kernel Name' in r:
        hdr = r; body = rows[i + 1:]; break
kn, mn, mv, mu, idc = (hdr.index(x) for x in ('Kernel Name', 'Metric Name', 'Metric Value', 'Metric Unit', 'ID'))
L = {}
for r in body:
    d = L.setdefault(int(r[idc]), {'name': r[kn].split('(')[0].replace('void ', '')[:46]})
    d[r[mn]] = (float(r[mv].replace(',', '')), r[mu])
ids = sorted(L)
start = max(i for i in ids if anchor in L[i]['name'])
tot = {}
for i in ids:
    if i < start: continue
    d = L[i]; t = d['gpu__time_duration.sum']
    ms = t[0] / 1e6 if t[1] == 'ns' else (t[0] / 1e3 if t[1] == 'us' else t[0])
    extra = ' '.join(f"{k.split('.')[0].replace('sm__', '').replace('smsp__', '')[:24]}={v[0]:.1f}" for k, v in d.items()
                     if k not in ('name', 'gpu__time_duration.sum'))
    tot[d['name']] = tot.get(d['name'], 0) + ms
    print(f"{d['name']:48s} {ms:8.3f} ms  {extra}")
print({k: round(v, 3) for k, v in tot.items()}, round(sum(tot.values()), 3))
