"""Summarise an ncu report into profiles/: key raw metrics (JSON) + per-source-line table (txt).
   python scripts/summarise_ncu.py <report.ncu-rep> <tag> <warp-steps of the profiled launch> "<what>" """
import csv, io, json, os, subprocess, sys
rep, tag, warp_steps, what = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ROUND = os.environ.get("ROUND", "r02")                      # file-name prefix under profiles/
WRITE_MIX = os.environ.get("INSTR_MIX", "1") == "1"         # only the default bench workload's capture feeds bench.py
sys.path.insert(0, ROOT)
import bench                                                 # kernel_source_digest(): ties the mix to the kernel source it was captured from
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u = rows[0], rows[1]
out = {"what": what, "launches": []}
want = ("gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.avg", "sm__cycles_active.avg", "lts__t_bytes.sum", "l1tex__t_bytes.sum")
for v in rows[2:]:
    m = {"kernel": v[h.index("Kernel Name")] if "Kernel Name" in h else ""}
    for k in want:
        if k in h:
            m[k] = [v[h.index(k)], u[h.index(k)]]
    for i, k in enumerate(h):
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k:
            m["stall_" + k.split("issue_stalled_")[1].replace("_per_issue_active.ratio", "")] = v[i]
    if "smsp__inst_executed.sum" in m:
        tot = float(m["smsp__inst_executed.sum"][0]) / warp_steps
        m["warp_instructions_per_warp_step"] = tot
        # FP64-pipe warp instructions: the pipe takes one per two cycles per SMSP, i.e. 2 per cycle per SM
        n_sm = 148
        f = float(m["sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"][0]) / 100 * 2 * float(m["sm__cycles_active.avg"][0]) * n_sm / warp_steps
        cyc = float(m["sm__cycles_elapsed.avg"][0]) * n_sm * 4 / warp_steps
        m["issue_model"] = {"fp64_inst_per_warp_step": f, "other_inst_per_warp_step": tot - f,
                            "issue_slots_per_warp_step": tot + f, "smsp_cycles_per_warp_step": cyc,
                            "issue_slot_frac": (tot + f) / cyc, "fp64_pipe_frac": 2 * f / cyc,
                            "note": "an FP64 warp instruction holds the SMSP issue port for 2 cycles on B200 (profiles/fp64_issue_probe_r01.txt): slots = 2*fp64 + other"}
        def mb(k):
            v, unit = m[k]
            return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
        if WRITE_MIX:
            json.dump({"source": os.path.basename(rep), "tag": tag, "what": what, "source_digest": bench.kernel_source_digest(),
                       "dram_bytes_per_launch": mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum"), **m["issue_model"]},
                      open(os.path.join(ROOT, "profiles", f"instr_mix_{ROUND}.json"), "w"), indent=1)
    out["launches"].append(m)
json.dump(out, open(os.path.join(ROOT, "profiles", f"ncu_{ROUND}_{tag}_summary.json"), "w"), indent=1)
lines = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_lines.py"), rep, str(warp_steps), "70"],
                       capture_output=True, text=True).stdout
open(os.path.join(ROOT, "profiles", f"ncu_{ROUND}_{tag}_source_lines.txt"), "w").write(what + "\n" + lines)
print(json.dumps(out["launches"][0], indent=1)[:1500])
