"""bench.py's reference arm (the CPU restatement timed on the host cores) runs without a GPU: one JSON line with the
contract's keys on rank 0, nothing and exit 0 on the other ranks."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                           "--workload", "1m-u"], capture_output=True, text=True, env=env, timeout=600)


def test_reference_arm_prints_the_contract_line():
    res = _run({})
    assert res.returncode == 0, res.stderr[-2000:]
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "particle-steps/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["gpu_launches"] == 0
    from oracle import wgsl_ref
    # the reference's WGSL transpiled to C++ when oracle/_ref is there (built where /root/reference exists), else the port
    assert line["cpu_baseline"]["kind"] == ("reference" if wgsl_ref.available() else "port")
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0)), "the CPU arm must use every host core"
    assert line["config"]["particles"] == 1_000_000 and line["config"]["same_config"] is True
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["config"]["workload"] == "1m-u"


def test_reference_arm_ignores_torchruns_single_thread_default():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the reference arm is an all-core CPU figure regardless."""
    res = _run({"OMP_NUM_THREADS": "1", "RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    assert res.returncode == 0, res.stderr[-2000:]
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))


def test_reference_arm_other_ranks_do_nothing():
    res = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert res.returncode == 0, res.stderr[-2000:]
    assert res.stdout.strip() == ""
