"""The scripts under examples/ run as written (copied to a temp dir with the round counts cut down)."""
import os
import re
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _stage(tmp_path, rel, **subs):
    dst = tmp_path / os.path.basename(rel)
    shutil.copytree(os.path.join(ROOT, "examples", rel), dst)
    cfg = dst / "fedml_config.yaml"
    text = cfg.read_text()
    for key, val in subs.items():
        text, n = re.subn(rf"(\n\s*{key}:\s*).*", rf"\g<1>{val}", text)
        assert n >= 1, key
    cfg.write_text(text)
    return dst


def _run(cwd, script, *extra, timeout=240):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    return subprocess.run([sys.executable, script, "--cf", "fedml_config.yaml", *extra], cwd=cwd, env=env, capture_output=True,
                          text=True, timeout=timeout)


@pytest.mark.parametrize("script", ["one_line.py", "step_by_step.py", "custom_data_and_model.py"])
def test_sp_examples(tmp_path, script):
    d = _stage(tmp_path, "simulation/sp_fedavg_mnist_lr", comm_round=2, frequency_of_the_test=1, client_num_in_total=20, client_num_per_round=4)
    r = _run(d, script)
    assert r.returncode == 0 and "'round': 1" in r.stdout + r.stderr, (r.stdout + r.stderr)[-1500:]


@pytest.mark.parametrize("example", ["mp_fedavg_mnist_lr", "mp_base_framework", "mp_decentralized_fl"])
def test_mp_examples_run_threaded(tmp_path, example):
    d = _stage(tmp_path, f"simulation/{example}", comm_round=2, client_num_in_total=8)
    r = _run(d, "main.py")
    assert r.returncode == 0, (r.stdout + r.stderr)[-1500:]


def test_centralized_example(tmp_path):
    d = _stage(tmp_path, "centralized", comm_round=1, epochs=1)
    r = _run(d, "main.py")
    assert r.returncode == 0, (r.stdout + r.stderr)[-1500:]


def test_cross_silo_grpc_example_three_processes(tmp_path):
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    d = _stage(tmp_path, "cross_silo/grpc_fedavg_mnist_lr", comm_round=2, grpc_base_port=port)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    procs = [subprocess.Popen([sys.executable, "server.py" if r == 0 else "client.py", "--cf", "fedml_config.yaml", "--rank", str(r),
                               "--run_id", "ex-grpc"], cwd=d, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in (0, 1, 2)]
    outs = []
    try:
        for p in procs:
            out, _ = p.communicate(timeout=240)
            outs.append(out)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    assert all(p.returncode == 0 for p in procs), "\n----\n".join(o[-800:] for o in outs)
    assert os.path.exists(d / "model_file_cache" / "global_model.pt")


def test_shipped_configs_are_in_the_documented_layout():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "annotate_configs.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout


def test_bench_cpu_config_prints_the_contract_line(tmp_path):
    """``bench.py --config sp_mnist_lr`` (BASELINE #1) runs without a GPU: one JSON line with the driver's keys, the SP simulator's
    number as the value and the client-batched engine next to it."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", "sp_mnist_lr", "--steps", "3", "--warmup", "1"],
                       cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-1500:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "dtype", "data", "config"):
        assert key in line, key
    assert line["metric"] == "client_updates_per_sec" and line["steps"] == 3 and line["value"] > 0 and line["n_gpus"] == 0
    assert line["client_batched"]["value"] > 0 and line["gpu_launches"] == 0
