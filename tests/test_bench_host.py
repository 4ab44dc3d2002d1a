"""Host-side logic of bench.py that needs no GPU: the per-kernel roofline table and the choice of the headline kernel."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_roofline_table_and_headline_kernel():
    import bench
    wl = dict(bench.WORKLOADS["c5_2m_800x800"], name="c5_2m_800x800")
    stats = {"R": 5_600_000, "pairs": 278_000_000}
    # (launches, total ms) over ONE step of 8 samples
    prof = {"gp_knn_kernel": (8, 8 * 0.41), "gp_render_bwd_kernel": (8, 8 * 0.40), "gp_rs_pass_kernel": (48, 48 * 0.043),
            "gp_preprocess_bwd_kernel": (8, 8 * 0.24)}
    roof, table, tot = bench.rooflines(prof, stats, wl, 1, 8, {"hbm_gbs": 6554.2}, 68.4, "test")
    # the search is the longest kernel but prunes (no algorithmic-work figure): the headline is the longest RATED kernel,
    # and the line says so
    assert roof["kernel"] == "gp_render_bwd_kernel" and "gp_knn_kernel" in roof["note"]
    assert table["gp_knn_kernel"].get("frac") is None
    flops = 45.0 * stats["pairs"]
    assert abs(roof["achieved"] - flops / 0.40e-3 / 1e12) < 1e-2 and abs(roof["frac"] - roof["achieved"] / 68.4) < 1e-3
    assert roof["traffic"] == bench.NCU_DRAM_BYTES["c5_2m_800x800"]["gp_render_bwd_kernel"]
    # HBM kernels: bytes per launch / time against the measured HBM peak; the sort: 4 depth passes over P + 2 over R
    pb = table["gp_preprocess_bwd_kernel"]
    assert pb["unit"] == "GB/s" and abs(pb["achieved"] - 535.0 * wl["N"] / 0.24e-3 / 1e9) < 1.0
    rs = table["gp_rs_pass_kernel"]
    want = (4 * 16.0 * wl["N"] + 2 * 16.0 * stats["R"]) / 6 / 0.043e-3 / 1e9
    assert abs(rs["achieved"] - want) < 1.0 and rs["launches_per_step"] == 48.0
    assert list(tot)[0] == "gp_knn_kernel"                       # per-step totals come sorted, longest first
    # when the longest kernel has a roofline there is no note
    prof["gp_render_bwd_kernel"] = (8, 8 * 0.50)
    roof2, _, _ = bench.rooflines(prof, stats, wl, 1, 8, {"hbm_gbs": 6554.2}, 68.4, "test")
    assert roof2["kernel"] == "gp_render_bwd_kernel" and "note" not in roof2


def test_workloads_match_baseline_configs():
    import json
    import bench
    cfgs = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert bench.HEADLINE in bench.WORKLOADS and bench.SAMPLES_PER_STEP == 8
    w = bench.WORKLOADS[bench.HEADLINE]
    assert w["N"] == 2_000_000 and (w["W"], w["H"]) == (800, 800) and w["Kp"] == 512 and w["K"] == 6
    assert "north_star" in cfgs and "configs" in cfgs
