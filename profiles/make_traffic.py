"""profiles/traffic.json from an `ncu --set full` capture of bench.py (read here, on the CPU box, with `ncu -i`).

usage: python profiles/make_traffic.py <workload> <capture.ncu-rep> <bench line .json of the same workload> [--csv out.csv]

The capture holds the k_trace / k_coop launches of a few frames run as ONE tile group (RTP_GROUPS=1: whole-frame
launches).  One frame = the rows from a camera-level closest-hit launch (FrameClosest<1>) up to the next one; ray
batches (C5): one RaysAny / RaysClosest launch.  Per kind (closest / anyhit, k_coop counted with the k_trace it
finishes) the frame's launches are summed: DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum), warp
instructions (smsp__inst_executed.sum), thread instructions (-> live lanes per instruction), duration under ncu.
`rays` comes from the bench line (roofline.rays_by_kind), so that bench.py can scale the counters to its launch."""
import csv
import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
COLS = dict(us="gpu__time_duration.sum", rd="dram__bytes_read.sum", wr="dram__bytes_write.sum", inst="smsp__inst_executed.sum",
            tinst="smsp__thread_inst_executed.sum", lanes="smsp__thread_inst_executed_per_inst_executed.ratio",
            l1="l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", l2="lts__t_sectors.sum",
            issue="smsp__issue_active.avg.pct_of_peak_sustained_active", warps="sm__warps_active.avg.pct_of_peak_sustained_active")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def kind_of(name):
    if "k_camera" in name or "k_coop_camera" in name:
        return "camera"
    if "k_paths" in name:
        return "paths"
    if "Closest" in name:
        return "closest"
    if "Any" in name:
        return "anyhit"
    return None


def main():
    workload, rep, bench = sys.argv[1], sys.argv[2], sys.argv[3]
    if rep.endswith(".csv"):  # `ncu -i capture.ncu-rep --page raw --csv` written on the GPU box (a .ncu-rep of 50 launches is too big to bring back)
        out = open(rep).read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r and not r[0].startswith("==")]
    H, U, data = rows[0], rows[1], rows[2:]
    ki = H.index("Kernel Name")

    def val(r, key):
        if COLS[key] not in H:
            return 0.0
        i = H.index(COLS[key])
        v = float(r[i].replace(",", "") or 0)
        return v * UNIT.get(U[i], 1.0)

    names = [r[ki] for r in data]
    cam = [i for i, n in enumerate(names) if ("FrameClosest<1>" in n and "k_trace" in n) or "k_camera" in n]
    if len(cam) >= 2:
        lo, hi = cam[0], cam[1]
    elif cam:
        lo, hi = cam[0], len(data)
    else:  # ray batches: the first launch alone
        lo, hi = 0, 1
    line = json.loads(open(bench).read().strip().splitlines()[-1])
    roof = line["roofline"]
    rays = roof.get("rays_by_kind") or {}
    if "--rays" in sys.argv:  # closest=N,anyhit=M (bench lines older than roofline.rays_by_kind)
        rays = {kv.split("=")[0]: float(kv.split("=")[1]) for kv in sys.argv[sys.argv.index("--rays") + 1].split(",")}
    if not rays:  # C5 line: one kind
        rays = {kind_of(names[lo]): roof["rays_per_launch_set"]}
    acc = {}
    for r in data[lo:hi]:
        k = kind_of(r[ki])
        if k is None:
            continue
        a = acc.setdefault(k, dict(us=0.0, dram=0.0, inst=0.0, tinst=0.0, l1=0.0, l2=0.0, launches=0, issue_w=0.0, warps_w=0.0))
        a["us"] += val(r, "us")
        a["dram"] += val(r, "rd") + val(r, "wr")
        a["inst"] += val(r, "inst")
        a["tinst"] += val(r, "tinst") if COLS["tinst"] in H else val(r, "inst") * val(r, "lanes")
        a["l1"] += val(r, "l1") * 32
        a["l2"] += val(r, "l2") * 32
        a["issue_w"] += val(r, "issue") * val(r, "us")
        a["warps_w"] += val(r, "warps") * val(r, "us")
        a["launches"] += 1
    path = os.path.join(HERE, "traffic.json")
    try:
        T = json.load(open(path))
    except Exception:
        T = {}
    T = {k: v for k, v in T.items() if isinstance(v, dict) and all(isinstance(x, dict) for x in v.values())}
    T["_how"] = ("per kind of trace launch (k_coop counted with its k_trace), summed over the launches of ONE frame run as one tile "
                 "group: dram_bytes = dram__bytes_read.sum + dram__bytes_write.sum, warp_inst = smsp__inst_executed.sum, lanes = "
                 "thread instructions / warp instructions, ncu_us = gpu__time_duration.sum (cold, serialised - not a bench "
                 "number), rays = the frame's rays of that kind (bench line).  Written by profiles/make_traffic.py from the "
                 "capture named in `from`; bench.py scales warp_inst and dram_bytes by rays(launch) / rays(entry).")
    ent = T.setdefault(workload, {})
    knames = {"closest": "k_trace<closest>", "anyhit": "k_trace<anyhit>", "camera": "k_camera", "paths": "k_paths"}
    if not cam:
        knames = {"closest": "k_trace<RaysClosest>", "anyhit": "k_trace<RaysAny>"}
    for k, a in acc.items():
        n = float(rays.get(k) or 0)
        ent[knames[k]] = dict(dram_bytes=a["dram"], warp_inst=a["inst"], lanes=round(a["tinst"] / a["inst"], 3), rays=n,
                              launches=a["launches"], ncu_us=round(a["us"], 2),
                              l1_bytes_per_ray=round(a["l1"] / n, 1) if n else None, l2_bytes_per_ray=round(a["l2"] / n, 1) if n else None,
                              dram_bytes_per_ray=round(a["dram"] / n, 1) if n else None,
                              warp_inst_per_ray=round(a["inst"] / n, 2) if n else None,
                              issue_active_pct=round(a["issue_w"] / a["us"], 1), warps_active_pct=round(a["warps_w"] / a["us"], 1),
                              **{"from": "profiles/" + os.path.basename(sys.argv[sys.argv.index("--csv") + 1]) if "--csv" in sys.argv else rep})
    json.dump(T, open(path, "w"), indent=1)
    if "--csv" in sys.argv:  # the judged columns of every captured launch, frame window marked
        dst = sys.argv[sys.argv.index("--csv") + 1]
        with open(dst, "w") as f:
            keys = [c for c in COLS.values() if c in H] + ["launch__registers_per_thread", "launch__grid_size",
                                                           "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
                                                           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
            keys = [k for k in keys if k in H]
            f.write("row,in_frame_window,kernel," + ",".join(f"{k} [{U[H.index(k)]}]" for k in keys) + "\n")
            for i, r in enumerate(data):
                short = r[ki].replace("rtp::", "").replace("<unnamed>::", "<").replace("unnamed>::", "").split("(")[0].replace("void ", "")
                f.write(f"{i},{int(lo <= i < hi)},\"{short}\"," + ",".join(r[H.index(k)].replace(",", "") for k in keys) + "\n")
    print(json.dumps(ent, indent=1))


if __name__ == "__main__":
    main()
