"""Experiment helper: aggregate an ncu launch list (--metrics gpu__time_duration.sum[,dram__bytes_*] --csv)
per kernel: launches, summed time, share, DRAM bytes per launch.  usage: launch_summary.py in.csv out.json"""
import collections, csv, json, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr, data = rows[0], rows[1:]
ix = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: collections.defaultdict(float)); cnt = collections.Counter()
for r in data:
    name = re.sub(r'^void (<unnamed>|\(anonymous namespace\))::', '', r[ix['Kernel Name']]); name = re.sub(r'\(.*', '', name)
    m = r[ix['Metric Name']]; v = float(r[ix['Metric Value']].replace(',', '')); u = r[ix['Metric Unit']]
    if m.startswith('dram__bytes'): v *= {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
    if m == 'gpu__time_duration.sum':
        v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1, 's': 1e3}.get(u, 1e-6); cnt[name] += 1
    agg[name][m] += v
tot = sum(a['gpu__time_duration.sum'] for a in agg.values())
out = {"_total_ms_serialised_cold": round(tot, 3)}
for name, a in sorted(agg.items(), key=lambda t: -t[1]['gpu__time_duration.sum']):
    n = cnt[name]; t = a['gpu__time_duration.sum']; rd = a['dram__bytes_read.sum']; wr = a['dram__bytes_write.sum']
    if t / tot > 0.002:
        print(f"{name[:60]:60s} launches={n:4d} ms={t:9.3f} share={t/tot*100:5.1f}% dram/launch={(rd+wr)/n/1e6:9.1f} MB  dram GB/s={(rd+wr)/1e6/t if t else 0:7.1f}")
    out[name] = {"launches": n, "ms": round(t, 3), "share_pct": round(t / tot * 100, 2), "dram_bytes_per_launch": int((rd + wr) / n)}
json.dump(out, open(sys.argv[2], 'w'), indent=1)
