"""bench.py's output contract (the driver parses ONE JSON line from stdout): keys, arms, single line.
CPU part: the reference arm on the smallest config.  GPU part: the B200 arm on the smallest config."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config", "e2e", "cpu_baseline"}


def _run(*args):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, "stdout must be exactly one line, got %d: %r" % (len(lines), p.stdout[:500])
    return json.loads(lines[0])


def test_reference_arm_prints_one_json_line():
    """--impl reference: the reference's own CPU code (oracle/_ref, else the C port) on a bounded sample, same metric / unit /
    config as the B200 arm, e2e = the line's own value with zero copy bytes."""
    d = _run("--impl", "reference", "--config", "c1", "--steps", "1", "--warmup", "0")
    assert d["impl"] == "reference" and BASE_KEYS <= set(d)
    assert d["metric"] == "Mpath-segments/s" and d["unit"] == "M segments/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["n_gpus"] == 1 and d["vs_baseline"] is None and d["dtype"] == "f32"
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_b200_arm_fails_loudly_without_a_gpu():
    """No CPU fallback: without a CUDA device the B200 arm exits non-zero instead of printing a number."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", "c1", "--steps", "1", "--warmup", "3", "--no-cpu"],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode != 0 and not p.stdout.strip()


@pytest.mark.gpu
def test_b200_arm_prints_one_json_line():
    d = _run("--config", "c1", "--steps", "2", "--warmup", "3")
    assert BASE_KEYS | {"roofline", "gpu_launches", "clocks"} <= set(d) and "impl" not in d
    assert d["metric"] == "Mpath-segments/s" and d["value"] > 0 and d["gpu_launches"] > 0
    r = d["roofline"]
    # the roof follows where the tree lives: shared memory -> un-fused fp32, beyond -> random sectors
    assert r["bound"] in ("fp32_unfused", "hbm_random_sector") and r["unit"] in ("TFLOP/s", "G sectors/s") and r["peak"] > 0
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3 and r["frac"] <= 1.0
    assert "traffic" in r and "algorithmic_gbs" in r and d["frame_crc"] and d["sum_total"] > 0
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert d["cpu_baseline"]["value"] > 0 and d["clocks"]["reasons"] is not None


@pytest.mark.gpu
def test_group_backend_prints_the_same_frame():
    """bench.py --backend group: the frame through srt_render_group equals the frame of the per-rank route (same Philox keys;
    on one GPU the same launches - CRC identical)."""
    a = _run("--config", "c1", "--steps", "1", "--warmup", "3", "--no-cpu")
    b = _run("--config", "c1", "--steps", "1", "--warmup", "3", "--backend", "group", "--gpus", "1")
    assert BASE_KEYS | {"gpu_launches", "clocks", "backend"} <= set(b) and b["backend"] == "group"
    assert b["value"] > 0 and b["e2e"]["value"] > 0 and b["gpu_launches"] > 0
    assert b["segments_per_step"] == a["segments_per_step"]
    assert abs(b["sum_total"] - a["sum_total"]) <= 1e-5 * abs(a["sum_total"])
