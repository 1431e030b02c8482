"""Interval copies with the packed bytes cut into more, smaller chunks per SM (AREAL_B200_INTERVAL_CHUNKS_PER_SM):
does a narrower window of concurrently streamed memory change the achieved bandwidth?"""
import json, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for cps in sys.argv[1:] or ["8", "32", "128", "512", "2"]:
    code = "import sys; sys.path.insert(0, %r)\nimport bench_kernels as BK; out=[]; BK.bench_intervals(out, None)\n" % root
    env = dict(os.environ, AREAL_B200_INTERVAL_CHUNKS_PER_SM=cps)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root, env=env)
    for line in r.stdout.splitlines():
        if '"kernel"' in line and "set_" not in line:
            d = json.loads(line)
            print(cps, d["kernel"][:60], d.get("n_intervals"), d["ms"], d["frac_of_measured_peak"], flush=True)
    if r.returncode:
        print(cps, "FAILED", r.stderr[-300:])
