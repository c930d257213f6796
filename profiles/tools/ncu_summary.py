"""Text summaries of an .ncu-rep (read on the CPU box with `ncu -i`): the metrics DESIGN.md quotes, and the warp instructions /
stall samples per source line of the first profiled kernel.
  python profiles/tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/r02_x_ncu   ->  ..._summary.txt, ..._lines.txt"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

METRICS = """gpu__time_duration.sum dram__bytes_read.sum dram__bytes_write.sum launch__registers_per_thread launch__grid_size
launch__block_size launch__occupancy_limit_shared_mem launch__occupancy_limit_registers
sm__warps_active.avg.pct_of_peak_sustained_active sm__throughput.avg.pct_of_peak_sustained_elapsed
smsp__issue_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum smsp__thread_inst_executed_per_inst_executed.ratio
l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active l1tex__t_sector_hit_rate.pct lts__t_sector_hit_rate.pct
smsp__sass_inst_executed_op_local_ld.sum""".split()


def run(*args):
    return subprocess.run(["ncu", *args], check=True, capture_output=True, text=True).stdout


def summary(rep, out):
    rows = list(csv.reader(io.StringIO(run("-i", rep, "--page", "raw", "--csv"))))
    head, units, vals = rows[0], rows[1], rows[2]
    col = {h: k for k, h in enumerate(head)}
    with open(out, "w") as f:
        f.write("# kernel: %s\n" % vals[col["Kernel Name"]])
        names = [m for m in METRICS if m in col] + sorted(h for h in head if re.match(r"smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio$", h) and "not_issued" not in h)
        for m in names:
            f.write("%-88s %s %s\n" % (m, vals[col[m]], units[col[m]]))


def lines(rep, out, top=80):
    """Per CUDA source line (the rows of the cuda,sass view whose address column is '-')."""
    text = run("-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass")
    inst = defaultdict(float); samp = defaultdict(float); thr = defaultdict(float); src = {}
    path, col = "", None
    for r in csv.reader(io.StringIO(text)):
        if len(r) == 2 and r[0] == "File Path":
            path = r[1].split("/")[-1]; continue
        if r and r[0] == "Line No":
            col = {}
            for k, h in enumerate(r):
                col.setdefault(h, k)
            continue
        if col is None or len(r) < len(col) or r[col["Address"]] != "-":
            continue
        try:
            key = (path, int(r[0]))
            inst[key] += float(r[col["Instructions Executed"]] or 0)
            samp[key] += float(r[col["Warp Stall Sampling (All Samples)"]] or 0)
            thr[key] += float(r[col["Thread Instructions Executed"]] or 0)
            src[key] = r[1].strip()
        except (ValueError, KeyError):
            continue
    ti, ts = sum(inst.values()) or 1.0, sum(samp.values()) or 1.0
    with open(out, "w") as f:
        f.write("total warp instructions attributed to source lines: %d\n" % ti)
        for key in sorted(inst, key=lambda k: -inst[k])[:top]:
            if inst[key] <= 0:
                break
            f.write("%5.2f%% inst %5.2f%% samples  lanes=%4.1f  %s:%d  %s\n" % (100 * inst[key] / ti, 100 * samp[key] / ts, thr[key] / inst[key], key[0], key[1], src[key][:110]))


if __name__ == "__main__":
    rep, stem = sys.argv[1], sys.argv[2]
    summary(rep, stem + "_summary.txt")
    try:
        lines(rep, stem + "_lines.txt")
    except Exception as e:                      # the source page layout differs between ncu versions
        print("lines: " + repr(e))
