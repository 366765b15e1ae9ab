"""Experiment helper (not part of the product): run bench.py with several option sets and
print one summary line each.  usage: python tools_sweep.py WORKLOAD "k=v,k=v" "k=v" ..."""
import json, os, subprocess, sys
wl = sys.argv[1]
for spec in sys.argv[2:]:
    opts = []
    for kv in spec.split(","):
        if kv and kv != "base":
            opts += ["--opt", kv]
    res = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"), "--workload", wl, "--steps", "2", "--warmup", "1", "--no-cpu"] + opts,
                         capture_output=True, text=True)
    try:
        d = json.loads(res.stdout.strip().splitlines()[-1])
        ks = {k["name"]: round(k["ms"] / d["steps"], 1) for k in d["kernels"]}
        print(f"{wl} [{spec}] ms/step={d['ms_per_step']:.1f} q/s={d['queries_per_sec']:.0f} frac={d['roofline']['whole_step']['frac']} {ks}", flush=True)
    except Exception as e:
        print(f"{wl} [{spec}] FAILED {e} {res.stderr[-300:]}", flush=True)
