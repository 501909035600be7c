"""bench.py contract checks that need no GPU: the byte / flop accounting
and the --impl reference arm (the CPU restatement on the host cores) printing one JSON line with the agreed keys."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_algorithmic_bytes_match_the_survey():
    b = bench.algorithmic_bytes(1)
    assert round(b["corr_fwd"] / 1e6, 3) == 11.710           # read a, b 6.291; write out 5.419
    assert round(b["corr_bwd"] / 1e6, 3) == 18.002
    assert round(b["warp_fwd"] / 1e6, 3) == 6.291
    assert round(b["warp_bwd"] / 1e6, 3) == 12.583
    assert b["downsample"] == 384 * 512 * 2 * 4 + sum(h * w * 2 * 4 for h, w in bench.DS_SIZES)
    b4k = bench.algorithmic_bytes(1, bench.H4K, bench.W4K)
    assert round(b4k["corr_fwd"] / 1e6, 1) == 497.7 and round(b4k["corr_bwd"] / 1e6, 1) == 765.1
    f = bench.corr_flops(32)
    assert round((f["corr_fwd"] + f["corr_bwd"]) / 1e9, 1) == 66.6


def test_executed_mma_flops_cover_the_useful_ones():
    useful = bench.corr_flops(32)
    for prec, products in (("tf32", 1), ("tf32x3", 3)):
        m = bench.executed_mma_flops(32, prec)
        assert m["corr_fwd"] >= products * useful["corr_fwd"] and m["corr_bwd"] >= products * useful["corr_bwd"]
        assert m["corr_fwd"] <= 8 * useful["corr_fwd"] and m["corr_bwd"] <= 8 * useful["corr_bwd"]
    assert bench.executed_mma_flops(32, "fp32") is None


def test_reference_arm_prints_one_json_line():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--batch", "2"], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["metric"] == "image_pairs_per_s" and line["unit"] == "pairs/s"
    assert line["higher_is_better"] is True and line["value"] > 0 and line["gpu_launches"] == 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in line["config"]
