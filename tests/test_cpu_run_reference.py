"""``python -m idc_models_b200.run_reference <script>``: the reference's own scripts, byte for byte as they are under
/root/reference, executed on this framework through the import stand-ins (tensorflow, tensorflow_federated, phe,
nest_asyncio, keras.utils.np_utils, matplotlib.pyplot). Folders are tiny, so the scripts' hard-coded dataset sizes make
some of their splits empty — what is checked is that every call the scripts make exists and behaves (two-phase fit with
the ``initial_epoch = history.epoch[-1]`` quirk, ModelCheckpoint-free logging, the plot, the secure round loop)."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")
Image = pytest.importorskip("PIL.Image")
pytestmark = pytest.mark.skipif(not REF.exists(), reason="the reference checkout is not mounted on this box")


def _folder(root: Path, sub: str, n: int, size: int) -> None:
    rng = np.random.default_rng(0)
    for c in ("0", "1"):
        (root / sub / c).mkdir(parents=True)
    for i in range(n):
        c = i % 2
        a = (rng.random((size, size, 3)) * 110 + (90 if c else 30)).clip(0, 255).astype(np.uint8)
        Image.fromarray(a, "RGB").save(root / sub / str(c) / f"p{i:04d}.png")


def _run(script: str, *args: str, timeout: int = 900, **extra_env) -> str:
    env = dict(os.environ, PYTHONPATH=str(ROOT), IDCB_DEVICE="cpu", **extra_env)
    r = subprocess.run([sys.executable, "-m", "idc_models_b200.run_reference", str(REF / script), *args],
                       capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout


def test_unmodified_reference_transfer_script_runs(tmp_path):
    _folder(tmp_path, "data/IDC_regular_ps50_idx5/100", 48, 50)
    (tmp_path / "logs").mkdir()
    out = _run("dist_model_tf_mobile.py", str(tmp_path))
    assert "Pre-training with 1 devices took" in out and "Fine-tuning with 1 devices took" in out
    assert "Number of layers in the base model:  155" in out
    hist = [l for l in out.splitlines() if l.startswith("{'loss'")]
    assert len(hist) == 2
    fine = eval(hist[1])                                  # the script prints history_fine.history
    assert len(fine["loss"]) == 11                        # epochs 9..19: the upstream off-by-one is reproduced
    assert (tmp_path / "logs" / "plot_dev1.png").read_bytes()[:8] == b"\x89PNG\r\n\x1a\n"


def test_unmodified_reference_secure_script_runs(tmp_path):
    _folder(tmp_path, "data/balanced_IDC_30k", 200, 12)
    out = _run("secure_fed_model.py", str(tmp_path), "2", "0", IDCB_PAILLIER_BITS="256")
    assert out.count("Training for client 0 took") == 2 and out.count("Training for client 1 took") == 2
    assert "Secure fed model took" in out
    assert "binary_accuracy" in out and "auroc" in out    # the script's own py_func AUROC metric was accepted


def test_unmodified_reference_fedavg_script_runs_twice(tmp_path):
    """fed_model.py: first run pre-trains and writes <path>/pretrained/cp.ckpt (the script's ``sys.path.exists`` is
    given its intended meaning by the runner), the second run loads it; non-IID listing on the second run"""
    _folder(tmp_path, "data/balanced_IDC_30k", 32, 50)
    out = _run("fed_model.py", str(tmp_path), "1", "iid", timeout=1500)
    assert "Pre-training took" in out and "saving model to" in out and "Federated training took" in out
    assert (tmp_path / "pretrained" / "cp.ckpt").exists()
    out = _run("fed_model.py", str(tmp_path), "2", "noniid", timeout=1500)
    assert "Loading pretrained model" in out and "Pre-training took" not in out
    rounds = [l for l in out.splitlines() if l.strip().startswith(("0,", "1,"))]
    assert len(rounds) == 2 and all(len(l.split(",")) == 5 for l in rounds)


def test_unmodified_reference_secure_script_encrypted_path(tmp_path):
    """percent = 0.2: the first weight tensor of every client is Paillier-encrypted scalar by scalar, averaged under
    encryption by the script's Server.aggregate and decrypted by its clients (short key for the test; the ragged
    ``np.array`` calls of the script get the object-array result of the NumPy it was written for)"""
    _folder(tmp_path, "data/balanced_IDC_30k", 120, 12)
    out = _run("secure_fed_model.py", str(tmp_path), "1", "0.2", IDCB_PAILLIER_BITS="256")
    for c in (0, 1):
        assert f"Encryption for client {c} took" in out and f"Decryption for client {c} took" in out
    assert "Secure fed model took" in out


def test_unmodified_reference_transfer_script_as_two_ranks(tmp_path):
    """launched with torchrun, the script's argument-less ``tf.distribute.MirroredStrategy()`` joins the job: two
    data-parallel CPU replicas (gloo), ``num_replicas_in_sync`` = 2 in the script's own Timer lines"""
    import socket
    _folder(tmp_path, "data/IDC_regular_ps50_idx5/100", 48, 50)
    (tmp_path / "logs").mkdir()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, PYTHONPATH=str(ROOT), IDCB_DEVICE="cpu")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), "-m", "idc_models_b200.run_reference",
                        str(REF / "dist_model_tf_mobile.py"), str(tmp_path)],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("Pre-training with 2 devices took") == 2
    assert r.stdout.count("Fine-tuning with 2 devices took") == 2
    assert (tmp_path / "logs" / "plot_dev2.png").exists()
