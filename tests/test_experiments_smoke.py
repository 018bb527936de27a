"""Every training entry point runs end to end on the CPU (single process, tiny synthetic problem) and its loss is finite."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, cwd):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    r = subprocess.run([sys.executable] + args, capture_output=True, text=True, cwd=str(cwd), env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def _losses(out, pattern):
    vals = [float(v) for v in re.findall(pattern, out)]
    assert vals and all(v == v and abs(v) < 1e9 for v in vals), out[-1500:]
    return vals


@pytest.mark.timeout(700)
def test_ogb_gcn_trainer(tmp_path):
    out = _run([os.path.join(ROOT, "experiments", "OGB", "main.py"), "--backend", "nccl", "--dataset", "arxiv", "--epochs", "3",
                "--scale", "0.02", "--log_dir", str(tmp_path / "logs"), "--save", str(tmp_path / "best.pth")], tmp_path)
    _losses(out, r"loss: ([-0-9.e+]+)")
    logs = os.listdir(tmp_path / "logs")
    for suffix in ("training_loss.log", "validation_accuracy.log", "test_results.log", "training_times.log"):
        assert any(f.endswith(suffix) for f in logs), (suffix, logs)
    assert any(f.endswith("timing_report_world1.json") for f in logs) and os.path.exists(tmp_path / "best.pth")
    out = _run([os.path.join(ROOT, "experiments", "OGB", "main.py"), "--dataset", "arxiv", "--epochs", "1", "--scale", "0.02",
                "--log_dir", str(tmp_path / "logs2"), "--resume", str(tmp_path / "best.pth")], tmp_path)
    _losses(out, r"loss: ([-0-9.e+]+)")


@pytest.mark.timeout(700)
def test_ogb_lsc_rgat_trainer(tmp_path):
    out = _run([os.path.join(ROOT, "experiments", "OGB-LSC", "main.py"), "--comm_type", "nccl", "--dataset", "synthetic",
                "--num_papers", "256", "--num_authors", "64", "--num_institutions", "8", "--epochs", "2"], tmp_path)
    _losses(out, r"loss ([-0-9.e+]+)")
    assert "validation accuracy" in out


@pytest.mark.timeout(700)
def test_graphcast_trainer(tmp_path):
    out = _run([os.path.join(ROOT, "experiments", "GraphCast", "train_graphcast.py"), "--test_run", "--use_synthetic_data",
                "--backend", "single", "--mesh_level", "2", "--iters", "2", "--save_dir", str(tmp_path)], tmp_path)
    _losses(out, r"loss ([-0-9.e+]+)")
    assert any(f.startswith("model_") and f.endswith(".pth") for f in os.listdir(tmp_path))
