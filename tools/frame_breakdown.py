"""Per-frame kernel share from an ncu launch list (gpu__time_duration.sum CSV). Usage: frame_breakdown.py <csv> [frame]"""
import collections, csv, re, sys
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith('==')]
rows = list(csv.DictReader(lines))
seq = [(re.sub(r'\(.*', '', x['Kernel Name']).replace('void ', '').replace('isr::', '')[:44], x['Grid Size'], float(x['Metric Value'])) for x in rows]
idx = [i for i, s in enumerate(seq) if 'pack_recurrent' in s[0]]
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
frame = seq[idx[k]:idx[k + 1]]
tot = sum(s[2] for s in frame)
print("frame total us %.1f launches %d" % (tot / 1e3, len(frame)))
agg = collections.OrderedDict()
nconv = 0
for s in frame:
    key = s[0]
    if 'conv3x3' in key:
        nconv += 1
        key += ' #%02d' % nconv if nconv > 22 else (' preblock' if nconv <= 2 else ' trunk' + ('A' if nconv % 2 else 'B'))
    agg.setdefault(key, [0, 0]); agg[key][0] += s[2]; agg[key][1] += 1
for k, v in agg.items():
    print(f"{k:58s} n={v[1]:2d} {v[0]/1e3:9.1f} us ({v[0]/v[1]/1e3:7.1f} each) {100*v[0]/tot:5.1f}%")
