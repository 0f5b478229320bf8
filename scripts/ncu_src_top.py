"""Top stall sites of one kernel from `ncu -i X.ncu-rep --page source --csv --kernel-name regex:K`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]
data = []
for r in rows[hi + 1:]:
    if len(r) != len(hdr) or r[0] == 'Address':
        break  # next kernel instance
    data.append(r)
si, src, ie = hdr.index('# Samples'), hdr.index('Source'), hdr.index('Instructions Executed')
stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[si]) for r in data)
print('total samples', tot, 'instructions', len(data))
agg = {hdr[i]: sum(int(r[i]) for r in data) for i in stalls}
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
    print('  %-24s %8d %5.1f%%' % (k, v, 100 * v / tot))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
top = sorted(enumerate(data), key=lambda kv: -int(kv[1][si]))[:n]
for k, r in sorted(top):
    why = max(stalls, key=lambda i: int(r[i]))
    print('%5d %7s %5.1f%% exec=%9s %-16s %s' % (k, r[si], 100 * int(r[si]) / tot, r[ie], hdr[why][6:], r[src].strip()[:100]))
