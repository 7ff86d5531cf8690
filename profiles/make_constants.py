"""profiles/traffic.json and profiles/hw_counters.json from the committed ncu summaries.

    python profiles/make_constants.py r02d panda_64m:4194304 ur5e_1m:1048576 panda_boxy_64m:4194304 h1_16m:2097152

reads profiles/<tag>_check_<workload>_metrics.txt (summarize_ncu.py output: one block per captured launch,
`=== kernel` header then `metric value unit` lines) and writes, per workload, the means over the captured
launches of each kernel: DRAM bytes per launch (traffic.json) and issue / FMA-pipe / occupancy counters
(hw_counters.json).  `rows` is the batch size of the profiled command (collect.sh).  bench.py prints both
as profile constants next to its live numbers.
"""
import collections
import json
import pathlib
import sys

HERE = pathlib.Path(__file__).parent
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}


def blocks(path):
    cur = None
    for line in path.read_text().splitlines():
        if line.startswith("=== "):
            cur = {"kernel": line[4:]}
            yield cur
        elif cur is not None and line.strip():
            parts = line.split()
            try:
                val = float(parts[1])
            except (IndexError, ValueError):
                continue
            unit = parts[2] if len(parts) > 2 else ""
            cur[parts[0]] = val * SCALE.get(unit, 1.0)


def main(tag, specs):
    traffic = {"_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch (mean of the captured launches) from this round's "
                        "ncu --set full captures (profiles/<tag>_check_<workload>_metrics.txt, make_constants.py); bench.py scales "
                        "them to its rows per launch and reports the SUM of both kernels as roofline.traffic, labelled as a profile constant"}
    hw = {"_note": "ncu --set full, --clock-control none, one capture per workload (profiles/<tag>_check_<workload>_metrics.txt, "
                   "make_constants.py): percent of peak sustained while active; means over the captured launches"}
    for spec in specs:
        name, rows = spec.split(":")
        path = HERE / f"{tag}_check_{name}_metrics.txt"
        per = collections.defaultdict(list)
        for b in blocks(path):
            key = "gjk_queue_kernel" if "gjk_queue" in b["kernel"] else ("check_kernel" if "check_kernel" in b["kernel"] and "recheck" not in b["kernel"] else None)
            if key:
                per[key].append(b)
        mean = lambda bs, k: sum(b.get(k, 0.0) for b in bs) / max(len(bs), 1)
        traffic[name] = {"rows": int(rows), "source": f"profiles/{path.name}"}
        hw[name] = {"rows": int(rows)}
        for key, bs in per.items():
            traffic[name][f"{key}_bytes"] = mean(bs, "dram__bytes_read.sum") + mean(bs, "dram__bytes_write.sum")
            hw[name][key] = {
                "issue_active_pct": round(mean(bs, "smsp__issue_active.avg.pct_of_peak_sustained_active"), 1),
                "pipe_fma_cycles_active_pct": round(mean(bs, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"), 1),
                "warps_active_pct": round(mean(bs, "sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
                "threads_per_inst": round(mean(bs, "smsp__thread_inst_executed_per_inst_executed.ratio"), 1),
                "registers": int(mean(bs, "launch__registers_per_thread")),
                "block_size": int(mean(bs, "launch__block_size")),
                "duration_us_under_ncu": round(mean(bs, "gpu__time_duration.sum"), 1),
                "launches": len(bs),
            }
    (HERE / "traffic.json").write_text(json.dumps(traffic, indent=1))
    (HERE / "hw_counters.json").write_text(json.dumps(hw, indent=1))
    print(json.dumps(hw, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
