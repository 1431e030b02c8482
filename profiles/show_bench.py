import sys,json
d=json.loads(sys.stdin.read())
print("step %.3f host %.3f bwd %.3f fwd %.3f gae %.3f clocks %s" % (d["ms_per_step"], d["host_enqueue_ms_per_step"], d["roofline"]["ms"], d["roofline_fwd"]["ms"], d["roofline_gae"]["ms"], d["clocks"]))
print({k: {a: round(b, 3) for a, b in v.items()} for k, v in d.get("kernel_ms_spread", {}).items()})
