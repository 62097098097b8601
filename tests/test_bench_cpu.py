"""bench.py's reference arm on the CPU (no GPU needed): the line it prints, the config object shared with the GPU arm, and
that the CPU legs use every host core even when the launcher exported OMP_NUM_THREADS=1 (torch.distributed.run does)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_small_run_uses_all_cores():
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--m", "4096", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "sketch_gflops" and line["unit"] == "GFLOP/s" and line["higher_is_better"] is True
    assert line["n_gpus"] == 2 and line["config"]["n_gpus"] == 2 and line["config"]["m"] == 4096 and line["config"]["d"] == 2048
    assert line["value"] > 0 and line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    cores = len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["cores"] == cores and line["cpu_baseline"]["kind"] in ("reference", "port")
    sys.path.insert(0, ROOT)
    import bench
    assert line["config"] == bench.c2_config(4096, "C", 2)          # the GPU arm prints the same object


def test_other_ranks_of_the_reference_arm_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--m", "4096", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert r.returncode == 0 and r.stdout.strip() == ""
