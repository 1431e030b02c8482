"""A/B: single-pass fused actor step (profiles/probe_rowres_cfg.py) with the current library and with
alt/libareal_b200_old.so, alternating, in fresh processes on one box."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
old = os.path.join(root, "alt", "libareal_b200_old.so")
cfgs = sys.argv[1:] or ["4x3"]
for tag, lib in (("new", None), ("old", old), ("new", None), ("old", old)):
    code = "import sys; sys.path.insert(0, %r); import areal_b200._lib as L\n" % root
    if lib:
        code += "L.LIB_PATH = %r\n" % lib
    probe = os.path.join(root, "profiles", "probe_rowres_cfg.py")
    code += "sys.argv = ['probe'] + %r\n__file__ = %r\nexec(open(%r).read())\n" % (cfgs, probe, probe)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root)
    for line in r.stdout.splitlines():
        print(tag, line, flush=True)
    if r.returncode:
        print(tag, "FAILED", r.stderr[-400:])
