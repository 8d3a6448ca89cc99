import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui, gi = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit'), hdr.index('Grid Size')
agg = collections.defaultdict(lambda: [0, 0.0])
seq = []
for r in rows[1:]:
    try: v = float(r[vi].replace(',', ''))
    except ValueError: continue
    u = r[ui]
    v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
    n = r[ki].split('(')[0].replace('void ', '').replace('bist::', '').replace('<unnamed>::', '')[-58:]
    agg[n][0] += 1; agg[n][1] += v
    seq.append((n, r[gi], v))
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-60s n=%4d total=%9.1f us avg=%8.1f us share=%.3f" % (k, v[0], v[1], v[1] / v[0], v[1] / tot))
print("total us", tot)
if len(sys.argv) > 2:
    for s in seq[:int(sys.argv[2])]: print("%-60s %-16s %8.1f" % s)
