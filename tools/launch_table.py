"""Per (kernel, grid) table from an ncu --csv launch list with several metrics per launch."""
import csv, re, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1], errors='replace')) if len(r) > 5]
h = next(i for i, r in enumerate(rows) if 'Kernel Name' in r); H = rows[h]
kn, mn, mv, gs, idc = H.index('Kernel Name'), H.index('Metric Name'), H.index('Metric Value'), H.index('Grid Size'), H.index('ID')
d = collections.OrderedDict()
for r in rows[h + 1:]:
    e = d.setdefault(r[idc], {'name': re.sub(r'\(.*', '', r[kn]).replace('void ', ''), 'grid': r[gs]})
    try: e[r[mn]] = float(r[mv].replace(',', ''))
    except ValueError: pass
agg = collections.OrderedDict()
for e in d.values():
    a = agg.setdefault((e['name'], e['grid']), [0, 0.0, 0, 0.0])
    a[0] += 1; a[1] += e.get('gpu__time_duration.sum', 0) / 1000; a[2] = e.get('launch__registers_per_thread', 0); a[3] += e.get('smsp__inst_executed.sum', 0)
tot = sum(a[1] for a in agg.values())
print('total us %.1f launches %d' % (tot, len(d)))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print('%-28s %-14s n %3d us/launch %7.1f regs %3d Minst/launch %6.2f share %.3f' % (k[0], k[1], a[0], a[1] / a[0], a[2], a[3] / a[0] / 1e6, a[1] / tot))
