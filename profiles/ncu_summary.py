"""Summarise an .ncu-rep: key raw metrics + top stall lines from the source page.
usage: python profiles/ncu_summary.py file.ncu-rep [n_top_lines]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 14
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts.sum", "sm__cycles_elapsed.avg.per_second"]
for r in rows[2:]:
    print("=" * 100)
    for k in keys:
        for i, h in enumerate(hdr):
            if h == k:
                print(f"{k} [{units[i]}] = {r[i]}")
    st = [(float(r[i]), h) for i, h in enumerate(hdr) if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h and r[i]]
    st.sort(reverse=True)
    print("top stalls (warps per issue):", ", ".join(f"{h.split('issue_stalled_')[1].split('_per')[0]}={v:.2f}" for v, h in st[:7]))
def hottest(view):
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", view], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    hi = next(i for i, r in enumerate(srows) if "Source" in r and any("Sampling" in c for c in r))
    h = srows[hi]
    ci = h.index("Source"); si = next(i for i, c in enumerate(h) if "Sampling (All" in c)
    body = [r for r in srows[hi + 1:] if len(r) > si and r[si] not in ("", "0")]
    body.sort(key=lambda r: -float(r[si]))
    tot = sum(float(r[si]) for r in body) or 1
    print("-" * 100); print(f"hottest {view} lines (share of warp-stall samples):")
    for r in body[:ntop]:
        print(f"{100 * float(r[si]) / tot:5.1f}%  {r[ci].strip()[:120]}")


for view in ("cuda", "sass"):
    try:
        hottest(view)
        break
    except Exception as e:
        print(f"source page ({view}) unavailable:", repr(e)[:200])
