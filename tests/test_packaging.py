"""Packaging assets: setup.py metadata, the files a wheel must carry, container / devcontainer / docs entry points
(reference: setup.py, docker/, .devcontainer/, docs/)."""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_setup_metadata():
    out = subprocess.run([sys.executable, "setup.py", "--name", "--version"], cwd=ROOT, capture_output=True, text=True,
                         timeout=120)
    assert out.returncode == 0, out.stderr
    lines = [l.strip() for l in out.stdout.splitlines() if l.strip()]
    assert lines[-2] == "metisfl_b200" and re.fullmatch(r"\d+\.\d+\.\d+", lines[-1])


def test_wheel_payload_is_declared():
    src = open(os.path.join(ROOT, "setup.py")).read()
    for pattern in ("lib/*.so", "_core*.so", "bin/*", "resources/**/*"):   # kernels, controller core, tools, certificates
        assert pattern in src
    assert '"metisfl"' in src                                              # the compatibility import name ships too
    assert os.path.isdir(os.path.join(ROOT, "metisfl_b200", "resources", "ssl_config", "default"))


def test_container_and_devcontainer_assets():
    d = open(os.path.join(ROOT, "docker", "Dockerfile")).read()
    assert "bdist_wheel" in d and "openssh-server" in d and "PasswordAuthentication no" in d
    e = open(os.path.join(ROOT, "docker", "entrypoint.sh")).read()
    assert "METISFL_SSHD" in e and os.access(os.path.join(ROOT, "docker", "entrypoint.sh"), os.X_OK)
    dc = json.load(open(os.path.join(ROOT, ".devcontainer", "devcontainer.json")))
    assert dc["build"]["dockerfile"] == "Dockerfile" and "pytest" in dc["postCreateCommand"]
    assert os.path.exists(os.path.join(ROOT, ".devcontainer", "Dockerfile"))


def test_docs_index_points_at_existing_pages():
    idx = open(os.path.join(ROOT, "docs", "README.md")).read()
    pages = re.findall(r"`([\w/]+\.md)`", idx)
    assert len(pages) >= 5
    for p in pages:
        path = os.path.normpath(os.path.join(ROOT, "docs", p))
        assert os.path.exists(path), p


def test_shipped_ckks_artefacts_encrypt_and_aggregate():
    """`resources/fheparams/cryptoparams/` (reference: `metisfl/resources/fheparams/cryptoparams/`): context + public
    key + eval-mult key and no private key - enough to encrypt and for the controller's weighted average."""
    import os

    import numpy as np

    import metisfl_b200
    from metisfl_b200 import _core
    from metisfl_b200.encryption import fhe
    d = os.path.join(os.path.dirname(metisfl_b200.__file__), "resources", "fheparams", "cryptoparams")
    assert sorted(os.listdir(d)) == ["cryptocontext.txt", "key-eval-mult.txt", "key-public.txt"]
    learner = fhe.CKKS(backend="host")
    learner.load_crypto_context_from_file(os.path.join(d, "cryptocontext.txt"))
    learner.load_public_key_from_file(os.path.join(d, "key-public.txt"))
    assert learner.batch_size == 4096 and learner.scaling_factor_bits == 52
    ct = learner.encrypt(np.linspace(-1, 1, 100))
    h = _core.ckks_header(ct)
    assert h["n_ct"] == 1 and h["n_values"] == 100 and len(ct) > 786_000
    controller = fhe.CKKS(backend="host")
    controller.load_crypto_context_from_file(os.path.join(d, "cryptocontext.txt"))
    avg = controller.compute_weighted_average([ct, ct], [0.5, 0.5])
    assert _core.ckks_header(avg)["n_values"] == 100
