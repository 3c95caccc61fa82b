"""Brief text summary of an .ncu-rep: per kernel the headline metrics, stall ratios and the most-stalled SASS lines.
    python scripts/ncu_brief.py gpurun_out/x.ncu-rep [kernel-regex] [top-n]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
kre = sys.argv[2] if len(sys.argv) > 2 else None
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__inst_executed.sum"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("====", d.get("Kernel Name", "?")[:110])
    for k in KEYS:
        if k in d:
            print(f"  {k:88s} {d[k]}")
    st = sorted(((float(v), k) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and v), reverse=True)
    print("  stalls/issue:", ", ".join(f"{k.split('stalled_')[1].split('_per_issue')[0]}={v:.2f}" for v, k in st[:7]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["--kernel-name", "regex:" + kre] if kre else []),
                     capture_output=True, text=True).stdout
blocks = src.split('"Kernel Name",')
for b in blocks[1:]:
    lines = list(csv.reader(io.StringIO(b)))
    print("---- source:", lines[0][0][:100])
    h = lines[1]
    i_src, i_s, i_ex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    data = [(int(r[i_s] or 0), r[i_src].strip(), int(r[i_ex] or 0), n) for n, r in enumerate(lines[2:]) if len(r) > i_s]
    tot = sum(x[0] for x in data) or 1
    print("   samples", tot, "sass lines", len(data))
    for s_, t_, ex, n in sorted(data, reverse=True)[:topn]:
        print(f"   {100 * s_ / tot:5.1f}%  #{n:5d} x{ex:8d}  {t_[:100]}")
