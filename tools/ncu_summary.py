#!/usr/bin/env python3
"""Turn an `ncu --set full` report of ONE ME launch sequence of bench.py into the JSON / markdown summaries kept under
profiles/ (bench.py reads `roofline.traffic` from the JSON by kernel name).

    ncu --set full --import-source on --clock-control none -k regex:"k_search16|k_search_blk|k_fp|k_post|k_hme_ctl" \\
        --launch-skip 12 -c 12 -o gpurun_out/full python bench.py --steps 1 --warmup 1 --concurrency 1 --no-cpu-baseline --no-extra-configs
    python tools/ncu_summary.py gpurun_out/full.ncu-rep profiles/r02_ncu_full_1080p_m4

Kernel names follow svtme_kernel_name(): the two k_search16 launches of a sequence are the pre-HME and the level-0
search in that order, k_search_blk<.,1> / <.,2> are levels 1 / 2."""
import csv
import io
import json
import subprocess
import sys

METRICS = {
    "gpu__time_duration.sum": "time_us", "launch__registers_per_thread": "registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write", "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "smsp__inst_executed.sum": "warp_instructions",
}
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(head)}
    n16 = 0
    launches, kernels = [], {}
    for r in rows[2:]:
        name = r[ix["Kernel Name"]]
        if "k_search16" in name:
            label = "k_search16 (pre-HME)" if n16 == 0 else "k_search16 (HME level 0)"
            n16 += 1
        elif "k_search_blk" in name:
            label = "k_search_blk (HME level 1)" if ", 1>" in name else "k_search_blk (HME level 2)"
        elif "k_fp_big" in name:
            label = "k_fp_big"
        elif "k_fp" in name:
            label = "k_fp"
        elif "k_post" in name:
            label = "k_post"
        else:
            label = "k_hme_ctl"
        rec = {"kernel": label}
        for m, key in METRICS.items():
            if m not in ix:
                continue
            v = float(r[ix[m]].replace(",", ""))
            u = units[ix[m]]
            rec[key] = v * UNIT.get(u, 1.0) if (key.startswith("dram") or key == "time_us") else v
        rec["dram_bytes"] = rec.get("dram_read", 0) + rec.get("dram_write", 0)
        launches.append(rec)
        k = kernels.setdefault(label, {"launches": 0, "time_us": 0.0, "dram_bytes": 0.0})
        k["launches"] += 1
        k["time_us"] += rec["time_us"]
        k["dram_bytes"] += rec["dram_bytes"]
        for key in ("registers", "warps_active_pct", "alu_pipe_pct", "issue_active_pct", "l2_hit_pct"):
            k[key] = rec.get(key)
    hme = sum(v["dram_bytes"] for n, v in kernels.items() if n in ("k_hme_ctl", "k_search16 (pre-HME)", "k_search16 (HME level 0)",
                                                                    "k_search_blk (HME level 1)", "k_search_blk (HME level 2)"))
    doc = {"report": rep, "what": "one ME launch sequence = one mini-GOP of 32 pictures, 1920x1080 8-bit preset 4; ncu --set full, --clock-control none; "
                                   "times are cold-cache and serialised by the profiler", "kernels": kernels, "launches": launches,
           "hme_stage_dram_bytes": hme, "sequence_dram_bytes": sum(v["dram_bytes"] for v in kernels.values()),
           "sequence_time_us": sum(v["time_us"] for v in kernels.values())}
    with open(out + ".json", "w") as f:
        json.dump(doc, f, indent=1)
    with open(out + ".md", "w") as f:
        f.write("# ncu --set full of one ME sequence (mini-GOP of 32 pictures, 1080p preset 4)\n\n")
        f.write("| launch | time us | regs | warps active % | ALU pipe % | issue active % | DRAM read MB | DRAM write MB | L2 hit % | warp instr M |\n|---|---|---|---|---|---|---|---|---|---|\n")
        for r in launches:
            f.write("| %s | %.1f | %d | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f |\n" % (
                r["kernel"], r["time_us"], r.get("registers", 0), r.get("warps_active_pct", 0), r.get("alu_pipe_pct", 0), r.get("issue_active_pct", 0),
                r.get("dram_read", 0) / 1e6, r.get("dram_write", 0) / 1e6, r.get("l2_hit_pct", 0), r.get("warp_instructions", 0) / 1e6))
        f.write("\nHME-stage launches (k_hme_ctl x 5, k_search16 x 2, k_search_blk x 2): %.0f MB of DRAM traffic per mini-GOP; whole sequence %.0f MB, %.2f ms.\n"
                % (hme / 1e6, doc["sequence_dram_bytes"] / 1e6, doc["sequence_time_us"] / 1e3))
    print(out + ".json", out + ".md")


if __name__ == "__main__":
    main()
