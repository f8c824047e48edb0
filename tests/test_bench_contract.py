"""bench.py's reference arm (CPU; the driver runs `bench.py --impl reference` beside the GPU arm): one JSON line on
stdout with the contract's keys, nothing else on stdout, no product import, non-zero ranks silent."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, args=()):
    env = dict(os.environ)
    env.update(extra_env or {})
    code = ("import runpy, sys; sys.argv = ['bench.py', '--impl', 'reference', '--config', '1', '--steps', '1', '--warmup', '1'%s]; "
            "runpy.run_path('bench.py', run_name='__main__'); "
            "sys.stderr.write('LOADED_PRODUCT=%%d\\n' %% int(any(m.startswith('snn_for_object_detection_b200') for m in sys.modules)))"
            % "".join(", %r" % a for a in args))
    return subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)


def test_reference_arm_prints_one_json_line_and_never_imports_the_product():
    r = _run()
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[:500]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "frames/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["n_gpus"] == 1 and d["vs_baseline"] is None
    assert d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert "workload" in d["config"] and "model" not in d["config"]
    assert "LOADED_PRODUCT=0" in r.stderr          # the arm walks oracle/containers.py, not the product's module tree


def test_reference_arm_is_silent_on_other_ranks():
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, args=("--gpus", "2"))
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""
