"""A/B of the interval copies (bench_kernels.bench_intervals, no reference rows) with the current library and
with alt/libareal_b200_old.so, alternating, in fresh processes on one box."""
import json, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
old = os.path.join(root, "alt", "libareal_b200_old.so")
for tag, lib in (("new", None), ("old", old), ("new", None), ("old", old)):
    code = "import sys; sys.path.insert(0, %r); import areal_b200._lib as L\n" % root
    if lib:
        code += "L.LIB_PATH = %r\n" % lib
    code += "import bench_kernels as BK; out=[]; BK.bench_intervals(out, None)\n"
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root)
    for line in r.stdout.splitlines():
        if '"kernel"' in line:
            d = json.loads(line)
            print(tag, d["kernel"][:60], d.get("n_intervals"), d["ms"], d["frac_of_measured_peak"], flush=True)
    if r.returncode:
        print(tag, "FAILED", r.stderr[-300:])
