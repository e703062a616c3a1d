"""Summarise ncu captures from gpurun_out/ into profiles/ (tracked).  Usage:
   python tools/summarize_ncu.py launches <launches.csv> <out.md>
   python tools/summarize_ncu.py kernel <report.ncu-rep> <out.md>"""
import collections, csv, subprocess, sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def launches(src, out):
    rows = list(csv.reader(open(src)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    ci = {h: i for i, h in enumerate(rows[hi])}
    data = rows[hi + 1:]
    # the last closed-loop step: from its solve_prologue3_kernel launch (round 2: one C call per solve) to the end;
    # captures of the round-1 path (no prologue kernel) start after the second-to-last ADMM launch
    pro = [i for i, r in enumerate(data) if "solve_prologue" in r[ci["Kernel Name"]]]
    adm = [i for i, r in enumerate(data) if "admm" in r[ci["Kernel Name"]]]
    start = pro[-1] if pro else (adm[-2] + 1 if len(adm) > 1 else 0)
    agg, tot = collections.OrderedDict(), 0.0
    for r in data[start:]:
        nm = r[ci["Kernel Name"]].split("(")[0]
        v = float(r[ci["Metric Value"]]); u = r[ci["Metric Unit"]]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        agg.setdefault(nm, [0.0, 0]); agg[nm][0] += v; agg[nm][1] += 1; tot += v
    with open(out, "w") as f:
        f.write(f"# Launch list of the last closed-loop step ({src})\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares).\n\n")
        f.write("| kernel | launches | total us | share |\n|---|---|---|---|\n")
        for k, (v, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
            f.write(f"| `{k[-90:]}` | {c} | {v:.1f} | {100 * v / tot:.1f}% |\n")
        f.write(f"\nTotal {tot:.1f} us over {sum(c for _, c in agg.values())} launches.\n")


def kernel(rep, out):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write(f"# ncu --set full summary of {rep}\n\n")
        for vals in rows[2:]:
            d = dict(zip(hdr, vals)); un = dict(zip(hdr, units))
            f.write(f"## `{d.get('Kernel Name', '')[:100]}`  grid {d.get('Grid Size')} block {d.get('Block Size')}\n\n")
            f.write("| metric | value | unit |\n|---|---|---|\n")
            for k in KEYS:
                if k in d:
                    f.write(f"| {k} | {d[k]} | {un.get(k, '')} |\n")
            f.write("\n")


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](sys.argv[2], sys.argv[3])
