"""Turns an `ncu --set full` report (gpurun_out/*.ncu-rep) into the small JSON/markdown summaries kept under profiles/.

    python profiles/summarize_ncu.py gpurun_out/prof_k2_r1.ncu-rep wls_moments_d1_lab r1

Writes profiles/<round>_<key>_full.json (selected raw metrics per profiled launch) and updates
profiles/ncu_summary.json[key] with the per-launch DRAM traffic that bench.py reports as roofline.traffic.
"""
import csv
import io
import json
import os
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "lts__t_sector_hit_rate.pct",
    "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.sum", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
]
UNIT_SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1}


def main():
    rep, key, rnd = sys.argv[1], sys.argv[2], sys.argv[3]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    launches = []
    for r in data:
        rec = {"kernel": r[hdr.index("Kernel Name")]}
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                try:
                    rec[w] = {"value": float(r[i].replace(",", "")), "unit": units[i]}
                except ValueError:
                    pass
        # warp-stall breakdown (WarpStateStats / SourceCounters sections): sampled stall reasons, largest first
        stalls = {}
        for i, name in enumerate(hdr):
            if name.startswith("smsp__pcsamp_warps_issue_stalled_") and not name.endswith("_not_issued"):
                try:
                    stalls[name[len("smsp__pcsamp_warps_issue_stalled_"):]] = float(r[i].replace(",", ""))
                except ValueError:
                    pass
        tot = sum(stalls.values())
        if tot > 0:
            rec["warp_stall_share"] = {k: round(v / tot, 4) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1]) if v / tot >= 0.005}
        launches.append(rec)
    here = os.path.dirname(os.path.abspath(__file__))
    json.dump({"report": os.path.basename(rep), "launches": launches}, open(os.path.join(here, f"{rnd}_{key}_full.json"), "w"), indent=1)

    def si(rec, name):
        m = rec[name]
        return m["value"] * UNIT_SCALE.get(m["unit"], 1)

    if not all("dram__bytes_read.sum" in l for l in launches):  # a sections-only capture (stall reasons): no traffic to record
        print(json.dumps(launches[0], indent=1))
        return
    traffic = [si(l, "dram__bytes_read.sum") + si(l, "dram__bytes_write.sum") for l in launches]
    dur = [si(l, "gpu__time_duration.sum") for l in launches]
    summ_path = os.path.join(here, "ncu_summary.json")
    summ = json.load(open(summ_path)) if os.path.exists(summ_path) else {}
    summ[key] = {"round": rnd, "kernel": launches[0]["kernel"], "profiled_launches": len(launches),
                 "dram_bytes_per_launch": sum(traffic) / len(traffic), "duration_s_per_launch_under_ncu": sum(dur) / len(dur),
                 "registers_per_thread": launches[0].get("launch__registers_per_thread", {}).get("value"),
                 "grid": launches[0].get("launch__grid_size", {}).get("value")}
    json.dump(summ, open(summ_path, "w"), indent=1)
    print(json.dumps(summ[key], indent=1))


if __name__ == "__main__":
    main()
