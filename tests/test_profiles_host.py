"""The committed profile artefacts bench.py reads at run time parse and carry what it looks for (CPU only)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_traffic_file_names_the_stages_of_the_bench_line():
    with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
        tj = json.load(f)
    for stage in ("preprocess", "emit_keys", "sort", "ranges_pack", "render_fwd", "render_bwd", "preprocess_bwd"):
        assert tj["bytes_per_launch"][stage] > 0, stage
    assert "note" in tj


def test_issue_active_comes_from_the_committed_capture():
    import bench
    for k in ("render_bwd", "render_fwd"):
        ia = bench.issue_active(k)
        assert ia is not None and 0.0 < ia["pct"] <= 100.0 and "r02_ncu_full_raw.csv" in ia["source"]
    assert bench.issue_active("no_such_kernel") is None


def test_configs_file_has_every_workload_with_its_reference_arm():
    lines = [json.loads(l) for l in open(os.path.join(ROOT, "profiles", "r02_configs.jsonl")) if l.strip()]
    seen = {(l["workload"], l["n_gpus"], l["impl"]) for l in lines}
    for w in ("cfg1", "cfg2", "train", "cfg3", "cfg4", "cfg5"):
        assert (w, 1, "ours") in seen and (w, 1, "reference") in seen, w
    for w in ("cfg2", "train", "cfg3", "cfg4", "cfg5"):
        assert (w, 8, "ours") in seen, w
    for l in lines:
        if l["impl"] == "ours" and l["n_gpus"] > 1 and l["workload"] in ("cfg2", "cfg5"):
            assert l["exchange_check"]["ok"] and l["exchange_check"]["identical_across_ranks"]
        if l["impl"] == "ours" and l.get("parity") and l["parity"].get("checked", True):
            assert l["parity"]["ok"], (l["workload"], l["n_gpus"])
