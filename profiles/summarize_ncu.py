"""Turn an .ncu-rep capture into the text summaries committed under profiles/.

    python profiles/summarize_ncu.py gpurun_out/prof_check.ncu-rep profiles/r01_check_ur5e

    python profiles/summarize_ncu.py <rep> <out> gjk_queue      # source page of the first launch matching the name

writes <out>_metrics.txt (raw-page counters of every captured launch) and <out>_source.txt
(top source lines by warp-stall samples for the first captured launch; needs -lineinfo).
"""

import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "sm__cycles_elapsed.avg.per_second", "sm__cycles_active.avg",
]
STALLS = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio"
STALL_NAMES = ["short_scoreboard", "long_scoreboard", "barrier", "wait", "branch_resolving", "no_instruction",
               "math_pipe_throttle", "mio_throttle", "lg_throttle", "dispatch_stall", "not_selected", "selected"]


def ncu(rep, page, extra=()):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout


def main(rep, out, kernel=None):
    rows = list(csv.reader(io.StringIO(ncu(rep, "raw"))))
    hdr, units = rows[0], rows[1]
    with open(out + "_metrics.txt", "w") as f:
        for r in rows[2:]:
            f.write(f"=== {r[hdr.index('Kernel Name')]}\n")
            for k in KEYS + [STALLS % s for s in STALL_NAMES]:
                if k in hdr:
                    f.write(f"{k:86s} {r[hdr.index(k)]} {units[hdr.index(k)]}\n")
    sel = ("-k", "regex:" + kernel, "-c", "1") if kernel else ()
    rows = list(csv.reader(io.StringIO(ncu(rep, "source", ("--print-source", "cuda,sass", *sel)))))
    sections, cur = [], None
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            cur = {"file": r[1], "rows": [], "hdr": None}
            sections.append(cur)
        elif r and r[0] == "Line No" and cur is not None:
            cur["hdr"] = r
        elif cur and cur["hdr"] and not (len(r) >= 2 and r[0] == "Function Name"):
            cur["rows"].append(r)
    files = []
    first = []
    for s in sections:  # the first launch's sections: stop when a file repeats
        if s["file"] in files:
            break
        files.append(s["file"])
        first.append(s)
    samp, inst, tinst, noinst = (collections.Counter() for _ in range(4))
    src = {}
    num = lambda x: float(x) if x.replace(".", "", 1).isdigit() else 0.0  # noqa: E731
    for s in first:
        h = s["hdr"]
        iS, iI, iT, iN = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("stall_no_inst")
        fn = s["file"].split("/")[-1]
        for r in s["rows"]:
            if not r[0].isdigit():
                continue
            k = (fn, int(r[0]))
            samp[k] += num(r[iS]); inst[k] += num(r[iI]); tinst[k] += num(r[iT]); noinst[k] += num(r[iN])
            if r[1] and k not in src:
                src[k] = r[1]
    T, I = sum(samp.values()) or 1, sum(inst.values()) or 1
    with open(out + ("_source.txt" if not kernel else f"_source_{kernel}.txt"), "w") as f:
        f.write(f"total samples {T:.0f}, warp instructions {I:.0f}\n")
        f.write("file:line  samples%  inst%  threads/inst  no_inst% of samples  source\n")
        for k, v in samp.most_common(40):
            f.write(f"{k[0]}:{k[1]:<4d} {100*v/T:5.1f}% {100*inst[k]/I:5.1f}% {tinst[k]/max(inst[k],1):5.1f} {100*noinst[k]/max(v,1):4.0f}%  {src.get(k,'')[:100]}\n")


if __name__ == "__main__":
    main(*sys.argv[1:4])
