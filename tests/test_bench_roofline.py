"""The roofline of the committed bench lines must be recomputable from committed evidence:
profiles/r02lines/bench_*.json (the driver's command on a B200) -> profiles/traffic.json (per workload and kind of launch) ->
profiles/r02_ncu_*_launches.csv (the `ncu --set full` capture, one frame's launches marked).  No GPU needed."""
import csv
import glob
import importlib
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINES = sorted(glob.glob(os.path.join(ROOT, "profiles", "r02lines", "bench_*.json")))


def kind_of(name):  # profiles/make_traffic.py
    if "k_camera" in name or "k_coop_camera" in name:
        return "k_camera"
    if "k_paths" in name:
        return "k_paths"
    if "RaysAny" in name:
        return "k_trace<RaysAny>"
    if "Closest" in name:
        return "k_trace<closest>"
    if "Any" in name:
        return "k_trace<anyhit>"
    return None


def load(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


@pytest.mark.parametrize("path", LINES, ids=[os.path.basename(p)[6:-5] for p in LINES])
def test_line_roofline_follows_from_traffic_json_and_the_capture(path):
    workload = os.path.basename(path)[6:-5]
    line = load(path)
    r = line["roofline"]
    assert r["bound"] == "issue" and r["frac"] is not None and 0.0 < r["frac"] <= 1.0
    assert r["contract"]["bound"] == "hbm" and r["contract"]["alg_bytes_per_ray"] > 0
    assert 0.0 < r["dram_frac"] < 1.0 and 0.0 < r["lane_efficiency"] <= 1.0
    ent = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[workload][r["kernel"]]
    # the line from traffic.json: warp instructions scaled to the launch's rays / live duration / issue peak
    scale = r["rays_per_launch_set"] / ent["rays"]
    achieved = ent["warp_inst"] * scale / (r["kernel_ms"] / 1e3) / 1e9
    assert achieved == pytest.approx(r["achieved"], rel=1e-6)
    assert achieved / r["peak"] == pytest.approx(r["frac"], rel=1e-6)
    assert ent["dram_bytes"] * scale == pytest.approx(r["traffic"], rel=1e-6)
    assert ent["lanes"] / 32.0 == pytest.approx(r["lane_efficiency"], rel=1e-6)
    # traffic.json from the capture: the marked frame's launches of that kind, summed
    rows = list(csv.reader(open(os.path.join(ROOT, ent["from"]))))
    head, rows = rows[0], rows[1:]
    col = {h.split(" [")[0]: i for i, h in enumerate(head)}
    unit = {h.split(" [")[0]: h.split(" [")[1].rstrip("]") if " [" in h else "" for h in head}
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    inst = dram = 0.0
    n = 0
    for row in rows:
        if row[col["in_frame_window"]] != "1" or kind_of(row[col["kernel"]]) != r["kernel"]:
            continue
        inst += float(row[col["smsp__inst_executed.sum"]])
        dram += float(row[col["dram__bytes_read.sum"]]) * mult[unit["dram__bytes_read.sum"]]
        dram += float(row[col["dram__bytes_write.sum"]]) * mult[unit["dram__bytes_write.sum"]]
        n += 1
    assert n == ent["launches"] and n > 0
    assert inst == pytest.approx(ent["warp_inst"], rel=1e-9)
    assert dram == pytest.approx(ent["dram_bytes"], rel=1e-6)


def test_every_workload_has_a_committed_line_and_capture():
    assert {os.path.basename(p)[6:-5] for p in LINES} >= {"c1", "c2", "c3", "c4", "c5"}
    T = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    for w in ("c1", "c2", "c3", "c4", "c5"):
        for k, ent in T[w].items():
            assert os.path.exists(os.path.join(ROOT, ent["from"])), (w, k)
            assert ent["rays"] > 0 and ent["warp_inst"] > 0 and 0 < ent["lanes"] <= 32


def test_bench_roofline_dict_matches_the_line():
    """bench.roofline_dict on the line's own inputs reproduces the line (same code path the driver's run takes)."""
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    line = load(os.path.join(ROOT, "profiles", "r02lines", "bench_c3.json"))
    r = line["roofline"]
    got = bench.roofline_dict("c3", r["kernel"], r["kernel_ms"], r["rays_per_launch_set"], line["config"]["triangles"], 148,
                              line["clocks"], bench.load_peaks())
    assert got["frac"] == pytest.approx(r["frac"], rel=1e-6) and got["bound"] == "issue"
    assert got["contract"]["alg_bytes_per_ray"] == r["contract"]["alg_bytes_per_ray"] == 1360
