import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
hdr = rows[h]
si, so = hdr.index('# Samples'), hdr.index('Source')
ie = hdr.index('Instructions Executed')
data = [r for r in rows[h+1:] if len(r) > si and r[si].isdigit()]
tot = sum(int(r[si]) for r in data)
print('total samples', tot, 'instrs', len(data))
stall = [x for x in hdr if x.startswith('stall_') and 'Not' not in x]
agg = {}
for r in data:
    for x in stall:
        v = r[hdr.index(x)]
        if v.isdigit(): agg[x] = agg.get(x, 0) + int(v)
print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
for i, r in enumerate(data):
    r.append(i)
for r in sorted(data, key=lambda r: -int(r[si]))[:topn]:
    st = {x: int(r[hdr.index(x)]) for x in stall if r[hdr.index(x)].isdigit() and int(r[hdr.index(x)]) > 0}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:2]
    print("%5d #%4d %6s %-70s %s" % (int(r[si]), r[-1], r[ie], r[so][:70], top))
