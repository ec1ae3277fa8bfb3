"""Summarise an .ncu-rep (ncu --set full) into markdown: one column per profiled launch, the counters DESIGN.md §5 uses.
    python tools/ncu_summary.py gpurun_out/X.ncu-rep > profiles/X.md"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H, U, R = rows[0], rows[1], rows[2:]
M = [("gpu__time_duration.sum", "duration"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
     ("launch__registers_per_thread", "regs/thread"), ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
     ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy"),
     ("smsp__inst_executed.sum", "warp instructions"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue-slot utilisation"),
     ("sm__inst_executed.avg.per_cycle_elapsed", "IPC (per SM, elapsed)"),
     ("smsp__thread_inst_executed_per_inst_executed.ratio", "active lanes / warp instruction"),
     ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe active"),
     ("sass__inst_executed_global_loads", "global load instr"), ("sass__inst_executed_shared_loads", "shared load instr"),
     ("sass__inst_executed_shared_stores", "shared store instr"), ("sass__inst_executed_local_loads", "local load instr"),
     ("sass__inst_executed_local_stores", "local store instr"),
     ("l1tex__t_sector_hit_rate.pct", "L1 hit rate"), ("lts__t_sectors.sum", "L2 sectors (32 B)"), ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
     ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
     ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard / issue"),
     ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait / issue"),
     ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not_selected / issue"),
     ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe / issue"),
     ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "stall branch / issue"),
     ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard / issue")]
names = [r[H.index("Kernel Name")].split("(")[0].replace("<unnamed>::", "").replace("void ", "") for r in R]
print("| counter | " + " | ".join(names) + " |")
print("|---|" + "---|" * len(names))
for key, label in M:
    if key not in H: continue
    i = H.index(key)
    print(f"| {label} (`{key}`) [{U[i]}] | " + " | ".join(r[i] for r in R) + " |")

if len(sys.argv) > 3 and sys.argv[2] == "--json":
    # python tools/ncu_summary.py X.ncu-rep --json <workload> : merge per-kernel facts into profiles/r1_ncu_kernels.json
    import json, os, re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = os.path.join(root, "profiles", sys.argv[4] if len(sys.argv) > 4 else "r2_ncu_kernels.json")
    db = json.load(open(out)) if os.path.exists(out) else {}
    sys.path.insert(0, root)
    import bench                                             # the content hash of csrc/ the capture ran on (bench.py drops stale captures)
    db["csrc_hash"] = bench.csrc_hash()
    wl = db.setdefault(sys.argv[3], {})
    def val(r, key, scale=1.0):
        i = H.index(key); u = U[i]
        v = float(r[i].replace(",", ""))
        mult = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)
        return v * mult * scale
    alias = {"k_trace_shadow<1, 1, 0>": "k_shadow_primary", "k_trace_shadow<1, 0, 0>": "k_shadow_primary", "k_trace_shadow<0, 1, 0>": "k_shadow_queue",
             "k_trace_shadow<0, 0, 0>": "k_shadow_queue", "k_trace_primary<0, 0>": "k_trace_primary", "k_trace_queue<0>": "k_trace_queue",
             "k_trace_shadow<1, 1>": "k_shadow_primary", "k_trace_shadow<1, 0>": "k_shadow_primary", "k_trace_shadow<0, 1>": "k_shadow_queue",
             "k_trace_shadow<0, 0>": "k_shadow_queue", "k_trace_shadow<1>": "k_shadow_primary", "k_trace_shadow<0>": "k_shadow_queue"}
    # per kernel: the mean over its launches in the report (the two waves of a frame differ a lot); utilisation and lanes
    # weighted by duration / instructions
    acc = {}
    for nm, r in zip(names, R):
        nm = alias.get(nm.strip(), nm.strip())
        a = acc.setdefault(nm, {"n": 0, "dram": 0.0, "inst": 0.0, "ms": 0.0, "util_ms": 0.0, "lanes_inst": 0.0})
        ms, inst = val(r, "gpu__time_duration.sum"), val(r, "smsp__inst_executed.sum")
        a["n"] += 1; a["dram"] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum"); a["inst"] += inst; a["ms"] += ms
        a["util_ms"] += val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active") * ms
        a["lanes_inst"] += val(r, "smsp__thread_inst_executed_per_inst_executed.ratio") * inst
    for nm, a in acc.items():
        wl[nm] = {"dram_bytes_per_launch": a["dram"] / a["n"], "issue_slot_util_pct": a["util_ms"] / a["ms"],
                  "warp_instructions": a["inst"] / a["n"], "active_lanes": a["lanes_inst"] / a["inst"],
                  "duration_ms_under_ncu": a["ms"] / a["n"], "launches_averaged": a["n"],
                  "source": "profiles/" + os.path.basename(rep).replace(".ncu-rep", ".md")}
    json.dump(db, open(out, "w"), indent=1)
