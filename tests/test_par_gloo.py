"""CPU coverage of the N>1 path (no GPU): world-size-2 gloo runs of the halo-exchange / all-reduce protocol on the oracle's
per-part data, and of bench.py's launch contract for the reference arm under torchrun."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(n, script, *args, port=29631):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(port), script] + list(args)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)


def test_par_cg_over_gloo_world2():
    r = _torchrun(2, os.path.join(ROOT, "tests", "par_gloo_worker.py"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PAR_GLOO_OK" in r.stdout


def test_bench_reference_arm_under_torchrun_world2():
    r = _torchrun(2, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                  "--cpu-cells", "12", "--cg-iters", "2", port=29632)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                         # rank 0 alone prints
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["unit"] == "GB/s"


def test_8_parts_on_2_ranks_over_gloo():
    """Several sub-domains per process (BASELINE config 3 on fewer GPUs than parts): message ordering by (source part,
    destination part) and local copies between co-resident parts, on the oracle's data."""
    r = _torchrun(2, os.path.join(ROOT, "tests", "par_gloo_multi_worker.py"), port=29633)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PAR_GLOO_MULTI_OK" in r.stdout
