"""bench.py's guard rails that need no GPU: a library experiment / profiling switch in the environment (CBX_SKIP drops
kernels — the results are garbage) must stop the bench before it measures anything, unless the caller declares a tuning
run; cfg3's prompt-length fixture is present."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = {k: v for k, v in os.environ.items() if not k.startswith("CBX_")}
    env.update(env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, env=env, timeout=300)


def test_bench_refuses_experiment_knobs():
    for knob in ("CBX_SKIP", "CBX_MEGA", "CBX_GEMV_STREAM", "CBX_BLOCK_STOP"):
        r = _run({knob: "1"})
        assert r.returncode != 0 and "refusing to run" in r.stderr and knob in r.stderr, (knob, r.stderr[-300:])
        assert r.stdout.strip() == "", "no JSON line may be printed"
    r = _run({"CBX_SKIP": "2"}, "--config", "cfg3")
    assert r.returncode != 0 and "refusing to run" in r.stderr


def test_cfg3_prompt_length_fixture():
    d = json.load(open(os.path.join(ROOT, "tests", "golden", "prompt_lengths.json")))
    assert d["n_chunks"] == 154 == len(d["n_text"]) and min(d["n_text"]) == 17 and max(d["n_text"]) == 484
