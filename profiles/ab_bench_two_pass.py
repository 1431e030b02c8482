"""A/B of bench.py's two-pass and single-pass blocks with the current library and with alt/libareal_b200_old.so."""
import json, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
old = os.path.join(root, "alt", "libareal_b200_old.so")
for tag, lib in (("new", None), ("old", old), ("new", None), ("old", old)):
    code = "import sys; sys.path.insert(0, %r); import areal_b200._lib as L\n" % root
    if lib:
        code += "L.LIB_PATH = %r\n" % lib
    code += "sys.argv = ['bench.py', '--steps', '20', '--warmup', '5', '--skip-e2e', '--skip-cpu']\nimport runpy; runpy.run_path(%r, run_name='__main__')\n" % os.path.join(root, "bench.py")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print(tag, "single ms", round(d["ms_per_step"], 4), "two-pass ms", round(d["two_pass"]["ms_per_step"], 4),
              "fwd", round(d["two_pass"]["fwd"]["ms"], 4), "bwd", round(d["two_pass"]["bwd"]["ms"], 4), flush=True)
    except Exception as e:
        print(tag, "FAILED", repr(e), r.stderr[-300:])
