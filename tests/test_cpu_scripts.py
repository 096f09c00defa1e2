"""Smoke: each of the five same-named scripts runs a tiny configuration on CPU with synthetic data and
prints the reference's output line formats (SURVEY.md section 4 "Smoke")."""
import re
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _run(args, tmp_path):
    r = subprocess.run([sys.executable, *args], cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r.stdout


@pytest.mark.parametrize("script,extra,layers", [
    ("dist_model_tf_mobile.py", [], 155),
    ("dist_model_tf_vgg.py", [], 19),
    ("dist_model_tf_dense.py", ["--arch", "densenet121", "--batch-size", "16"], 427),
])
def test_transfer_scripts(script, extra, layers, tmp_path):
    out = _run([script, str(tmp_path), "--device", "cpu", "--dataset-size", "64", "--initial-epochs", "2",
                "--fine-tune-epochs", "1", "--steps-per-epoch", "1", *extra], tmp_path)
    assert re.search(r"Pre-training with 1 devices took [\d.e-]+ seconds", out)
    assert f"Number of layers in the base model:  {layers}" in out
    assert re.search(r"Fine-tuning with 1 devices took [\d.e-]+ seconds", out)
    assert out.count("'val_accuracy'") == 2
    assert (tmp_path / "logs" / "plot_dev1.png").read_bytes()[:8] == b"\x89PNG\r\n\x1a\n"


def test_fed_model_script_trains_then_resumes_from_checkpoint(tmp_path):
    args = ["fed_model.py", str(tmp_path), "2", "iid", "--device", "cpu", "--dataset-size", "200", "--clients", "5",
            "--pretrain-epochs", "1", "--steps-per-epoch", "2", "--arch", "mobilenet_v2", "--fine-tune-at", "140"]
    out = _run(args, tmp_path)
    assert "saving model to" in out and "Pre-training took" in out
    assert "Starting federated training" in out and "Initial model: " in out
    assert len(re.findall(r"^ ?\d+, [\d.]+, [\d.]+, [\d.]+, [\d.]+ $", out, flags=re.M)) == 2
    assert (tmp_path / "pretrained" / "cp.ckpt").exists()
    out2 = _run(args, tmp_path)
    assert "Loading pretrained model" in out2 and "Pre-training took" not in out2


def test_secure_fed_model_script(tmp_path):
    out = _run(["secure_fed_model.py", str(tmp_path), "2", "0.5", "--device", "cpu", "--dataset-size", "300",
                "--epochs", "1", "--verbose", "0"], tmp_path)
    assert out.count("Training for client 0 took") == 2 and out.count("Encryption for client 1 took") == 2
    assert out.count("Decryption for client 0 took") == 2
    assert len(re.findall(r"^[\d.]+ [\d.]+ [\d.]+$", out, flags=re.M)) == 2
    assert "Secure fed model took" in out


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` prints one JSON line {"impl": "reference", "unavailable": ...} and exits 0 (the
    reference needs TensorFlow 2.1 / TFF / phe, none installable offline); the product arm refuses to run without CUDA."""
    import json
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "3"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and isinstance(line["unavailable"], str) and line["unavailable"]
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([sys.executable, str(root / "bench.py")], capture_output=True, text=True, timeout=300)
        assert r.returncode == 2 and "CUDA" in r.stderr
