"""ncu capture -> the two files bench.py's roofline block reads and the judge can re-derive numbers from:

  profiles/<tag>_ncu_raw.csv   one row per captured launch, the metrics below in BASE units (a column subset of
                               `ncu -i <rep> --page raw --csv --print-units base`; the .ncu-rep itself stays out of git)
  profiles/<tag>_traffic.json  per kernel class (trace_rays_kernel / shade_kernel): DRAM and L2 bytes per live ray-bounce,
                               L1 wavefront %, issue-active, lanes per instruction ... + the build id of the librtcore.so
                               the capture ran (bench.py only uses it for a line of the SAME build)

usage (here, no GPU):  python tools/ncu_summary.py gpurun_out/prof.ncu-rep gpurun_out/ncu_target.json <tag> ["limiter sentence"]
`ncu_target.json` is written by tools/ncu_target.py in the plain (un-profiled) run of the same command: live rays per
bounce, build id, workload."""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = {  # short name -> regex on the ncu metric name
    "duration_ns": r"^gpu__time_duration\.sum$",
    "dram_read_bytes": r"^dram__bytes_read\.sum$",
    "dram_write_bytes": r"^dram__bytes_write\.sum$",
    "lts_sectors": r"^lts__t_sectors\.sum$",  # 32-byte sectors the L2 tag stage handled (reads + writes + atomics)
    "l1_wavefront_pct": r"^l1tex__data_pipe_lsu_wavefronts\.(avg|sum)\.pct_of_peak_sustained_elapsed$",
    "l1_wavefronts_shared": r"^l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$",
    "l1_global_ld_requests": r"^l1tex__t_requests_pipe_lsu_mem_global_op_ld\.sum$",
    "l1_global_ld_sectors": r"^l1tex__t_sectors_pipe_lsu_mem_global_op_ld\.sum$",
    "issue_active": r"^smsp__issue_active\.avg\.per_cycle_active$",
    "lanes_per_inst": r"^smsp__thread_inst_executed_per_inst_executed\.ratio$",
    "warp_inst": r"^smsp__inst_executed\.sum$",
    "l1_hit_pct": r"^l1tex__t_sector_hit_rate\.pct$",
    "l2_hit_pct": r"^lts__t_sector_hit_rate\.pct$",
    "registers": r"^launch__registers_per_thread$",
    "dram_pct": r"^gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed$",
    "warps_active_pct": r"^sm__warps_active\.avg\.pct_of_peak_sustained_active$",
}


def read_raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-units", "base"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, data = rows[head], rows[head + 2:]
    cols = {}
    for short, pat in WANT.items():
        for j, n in enumerate(names):
            if re.search(pat, n):
                cols[short] = j
                break
    k = names.index("Kernel Name")
    launches = []
    for r in data:
        if len(r) <= k or not r[k]:
            continue
        d = {"kernel": r[k]}
        for short, j in cols.items():
            try:
                d[short] = float(r[j].replace(",", ""))
            except (ValueError, IndexError):
                d[short] = None
        launches.append(d)
    return launches, {s: names[j] for s, j in cols.items()}


def main():
    rep, target_json, tag = sys.argv[1:4]
    limiter = sys.argv[4] if len(sys.argv) > 4 else None
    target = json.load(open(target_json))
    launches, metric_names = read_raw(rep)
    live = target["live_per_bounce"]
    per_class = {}
    counters = {"trace_rays_kernel": 0, "shade_kernel": 0}
    raw_path = os.path.join(ROOT, "profiles", f"{tag}_ncu_raw.csv")
    fields = ["kernel", "class", "bounce", "live_rays"] + list(WANT)
    with open(raw_path, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=fields)
        w.writeheader()
        for d in launches:
            cls = next((c for c in counters if c in d["kernel"]), None)
            if cls is None:
                continue
            b = counters[cls]
            counters[cls] += 1
            d["class"], d["bounce"], d["live_rays"] = cls, b, live[b] if b < len(live) else 0
            w.writerow({k: d.get(k) for k in fields})
            per_class.setdefault(cls, []).append(d)
    summary = {"build_id": target["build_id"], "kernel_id": target.get("kernel_id"), "workload": target["workload"], "raw_csv": os.path.relpath(raw_path, ROOT),
               "source": f"ncu --set full --clock-control none of `{target['command']}`; launches of bounces 0..{max(counters.values()) - 1} of one pass "
                         "(they hold the %.1f %% of the pass's live ray-bounces that the per-ray figures are taken over)" %
                         (100.0 * sum(live[:max(counters.values())]) / max(1, sum(live))),
               "metric_names": metric_names, "limiter": limiter}
    for cls, ds in per_class.items():
        rays = sum(d["live_rays"] for d in ds)
        t = sum(d["duration_ns"] for d in ds)

        def wavg(key):  # duration-weighted mean of a per-launch rate / percentage
            vals = [(d[key], d["duration_ns"]) for d in ds if d.get(key) is not None]
            return sum(v * w for v, w in vals) / max(1e-9, sum(w for _, w in vals)) if vals else None

        summary[cls] = {
            "launches": len(ds), "live_ray_bounces": rays, "duration_ms": t * 1e-6,
            "dram_bytes_per_live_ray_bounce": sum((d["dram_read_bytes"] or 0) + (d["dram_write_bytes"] or 0) for d in ds) / max(1, rays),
            "lts_bytes_per_live_ray_bounce": 32.0 * sum(d.get("lts_sectors") or 0 for d in ds) / max(1, rays),
            "l1_global_ld_sectors_per_live_ray_bounce": sum(d.get("l1_global_ld_sectors") or 0 for d in ds) / max(1, rays),
            "l1_global_ld_warp_requests_per_live_ray_bounce": sum(d.get("l1_global_ld_requests") or 0 for d in ds) / max(1, rays),
            "warp_inst_per_live_ray_bounce": sum(d["warp_inst"] or 0 for d in ds) / max(1, rays),
            "l1_wavefront_pct": wavg("l1_wavefront_pct"), "issue_active": wavg("issue_active"), "lanes_per_inst": wavg("lanes_per_inst"),
            "l1_hit_pct": wavg("l1_hit_pct"), "l2_hit_pct": wavg("l2_hit_pct"), "dram_pct_of_peak": wavg("dram_pct"),
            "registers": ds[0].get("registers"),
            "per_bounce": [{k: d.get(k) for k in ("bounce", "live_rays", "duration_ns", "lanes_per_inst", "issue_active", "l1_wavefront_pct", "l1_hit_pct",
                                                  "l2_hit_pct", "dram_read_bytes", "dram_write_bytes", "lts_sectors", "warp_inst", "l1_global_ld_sectors", "l1_global_ld_requests",
                                                  "l1_wavefronts_shared")} for d in ds],
        }
    out = os.path.join(ROOT, "profiles", f"{tag}_traffic.json")
    json.dump(summary, open(out, "w"), indent=1)
    print(json.dumps({k: (v if not isinstance(v, dict) or k not in per_class else {kk: vv for kk, vv in v.items() if kk != "per_bounce"}) for k, v in summary.items()}, indent=1))


if __name__ == "__main__":
    main()
