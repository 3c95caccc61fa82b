#!/usr/bin/env python
"""Markdown summary of an `ncu --set full` report of one bench step (scripts/profile_step.py):
per kernel duration, DRAM bytes, registers, grid, tensor-pipe activity, L1 data-pipe shares.
Usage: ncu_step_summary.py report.ncu-rep > profiles/rNN_ncu_step.md"""
import csv, io, subprocess, sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}


def val(r, k, scale=1.0):
    try:
        return float(r[ix[k]]) * scale
    except (KeyError, ValueError, IndexError):
        return float("nan")


def to_mb(r, k):
    u = units[ix[k]]
    f = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
    return val(r, k, f)


def to_us(r, k):
    u = units[ix[k]]
    f = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
    return val(r, k, f)


print(f"# ncu --set full, one bench step (configs[1], bf16x3): {sys.argv[1]}\n")
print("Per-launch numbers are cold-cache and serialised (ncu replays each kernel); compare shares.\n")
print("| # | kernel | us | DRAM rd MB | DRAM wr MB | regs | grid x block | tensor pipe % | L1 data pipe: LSU % | TC operand % |")
print("|---:|---|---:|---:|---:|---:|---|---:|---:|---:|")
tot = 0.0
agg = {}
for r in rows[2:]:
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "")
    us = to_us(r, "gpu__time_duration.sum")
    rd, wr = to_mb(r, "dram__bytes_read.sum"), to_mb(r, "dram__bytes_write.sum")
    tot += us
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += us; a[2] += rd; a[3] += wr
    print(f"| {r[ix['ID']]} | `{name[:60]}` | {us:.1f} | {rd:.1f} | {wr:.1f} | {r[ix['launch__registers_per_thread']]} | "
          f"{r[ix['launch__grid_size']]} x {r[ix['launch__block_size']]} | "
          f"{val(r, 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{val(r, 'l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed'):.1f} |")
print(f"\nTotal {tot:.1f} us over {len(rows) - 2} launches.\n")
print("| kernel | launches | total us | share | DRAM rd+wr MB per launch |")
print("|---|---:|---:|---:|---:|")
for name, (n, us, rd, wr) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name[:60]}` | {n} | {us:.1f} | {us / tot:.3f} | {(rd + wr) / n:.1f} |")
