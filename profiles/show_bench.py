import sys, json
d = json.loads(sys.stdin.read())
print("mode %s value %.3f M tok/s step %.3f ms host %.3f  clocks %s" % (d["mode"], d["value"] / 1e6, d["ms_per_step"], d["host_enqueue_ms_per_step"], d["clocks"]))
sp, tp = d["single_pass"], d["two_pass"]
print("single-pass: step %.3f ms, fused kernel %.3f ms = %.0f GB/s (%.3f of peak)" % (sp["ms_per_step"], sp["ms"], sp["achieved"], sp["frac"]))
print("two-pass:    step %.3f ms, fwd %.3f ms (%.3f), bwd %.3f ms (%.3f, active %.2f)" % (tp["ms_per_step"], tp["fwd"]["ms"], tp["fwd"]["frac"], tp["bwd"]["ms"], tp["bwd"]["frac"], tp["bwd"]["active_row_frac"]))
print({k: {a: round(b, 3) for a, b in v.items()} for k, v in d.get("kernel_ms_spread", {}).items()})
if d.get("e2e"): print("e2e", d["e2e"])
if d.get("cpu_baseline"): print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
