"""CPU checks of bench.py's contract pieces that need no GPU: the algorithmic byte counts of SURVEY §8(d)
(what roofline.achieved divides by the kernel time), the workload table, and the reference arm's JSON line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_algorithmic_bytes_match_survey_8d():
    # SpMV: nnz*(sizeof(T)+4) + (h+1)*4 + w*sizeof(T) + 2*h*sizeof(T); SpMM: the dense terms times k
    h = 2048 * 2048
    assert bench.algorithmic_bytes("spmv", h, h, 5 * h - 4 * 2048, 1, 8) == 369_000_452          # config 2
    h = 1 << 24
    assert bench.algorithmic_bytes("spmm", h, h, 1 << 29, 64, 8) == 32_279_363_588               # config 3
    h = 1 << 21
    assert bench.algorithmic_bytes("spmm", h, h, 1 << 27, 256, 8) == 14_503_903_236              # config 5
    # gather model of config 3 (every nnz fetches its own B row): 298.6 GB
    assert round(bench.gather_bytes("spmm", 1 << 24, 1 << 24, 1 << 29, 64, 8) / 1e9, 1) == 298.6


def test_default_workload_is_the_north_star_config():
    """BENCH / SCALE measure configs[2] (16M x 32 uniform, fp64 SpMM k=64, method 2) at every N."""
    assert bench.DEFAULT_WORKLOAD == "c3"
    cfgs = json.load(open(os.path.join(ROOT, "BASELINE.json")))["configs"]
    assert "16M rows" in cfgs[2] and "k=64" in cfgs[2] and "method 2" in cfgs[2]
    assert set(bench.OTHERS_N1) == {"c2", "c3v", "c4", "c5"} and set(bench.OTHERS_NX) == {"c4", "c5"}


def test_workload_table_names_the_baseline_configs():
    cfgs = json.load(open(os.path.join(ROOT, "BASELINE.json")))["configs"]
    wl = bench.WORKLOADS
    assert wl["c2"]["op"] == "spmv" and wl["c2"]["nx"] * wl["c2"]["ny"] == 4 * 1024 * 1024 and "Laplacian" in cfgs[1]
    assert wl["c3"]["rows"] == 1 << 24 and wl["c3"]["deg"] == 32 and wl["c3"]["k"] == 64 and wl["c3"]["method"] == 2
    assert wl["c4"]["dtype"] == "f32" and wl["c4"]["scale"] == 23 and "R-MAT" in cfgs[3]
    assert wl["c5"]["rows"] == 1 << 21 and wl["c5"]["bw"] == 64 and wl["c5"]["k"] == 256 and wl["c5"]["method"] == 1


def test_reference_arm_line_on_a_small_instance():
    """bench.py --impl reference runs without a GPU (host cores only) and prints the contract's keys."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "1", "--scale-down", "64"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["unit"] == "GFLOP/s" and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["value"] > 0 and d["gpu_launches"] == 0


def test_reference_arm_other_ranks_exit_quietly():
    """Under torchrun (N > 1) rank 0 alone runs and prints the reference arm; the other ranks exit 0 without work."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29555")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "1", "--scale-down", "64"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == "", r.stdout + r.stderr


def test_gpu_arm_fails_loudly_without_a_gpu():
    """No CPU fallback: without a CUDA device the product arm stops with a message, it does not compute on the host."""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode != 0 and "no CPU fallback" in (r.stdout + r.stderr)
