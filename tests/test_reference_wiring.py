"""the hook wired into the REFERENCE's own classes.  integration/ref_wiring_check.py runs
in a subprocess whose sys.path starts with the reference's `vllm` package -- baseline/_ref (the offline
`pip install --target` of the unmodified reference; it travels to the GPU box) or /root/reference (build
container only) -- because site-packages holds an unrelated vllm release.  Skipped where neither exists."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_root():
    for cand in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.exists(os.path.join(cand, "vllm", "attention", "trianglemix_attention.py")):
            return cand
    return None


def _run(extra):
    ref = _reference_root()
    if ref is None:
        pytest.skip("no reference vllm package (baseline/_ref or /root/reference)")
    env = dict(os.environ, PYTHONPATH=ref, PYTHONDONTWRITEBYTECODE="1", VLLM_LOGGING_LEVEL="ERROR")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "integration", "ref_wiring_check.py")] + extra,
                       env=env, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def test_reference_attention_layer_installs_the_backend():
    res = _run([])
    assert res["vllm_version"].startswith("0.11")                 # the reference's vllm, not site-packages'
    assert res["impl"] == ["TriangleMixImpl"] * 3 and res["modes"] == [0, 1, 1] and res["metadata"] == "ok"


@pytest.mark.gpu
def test_reference_attention_forward_runs_our_kernels():
    res = _run(["--gpu"])
    assert len(res["forward"]) == 3 and all(e[1] == 0 and e[2] == 0 for e in res["forward"]), res


def test_patch_file_is_present_and_names_the_hook_sites():
    with open(os.path.join(ROOT, "integration", "vllm_trianglemix.patch")) as f:
        patch = f.read()
    for must in ("vllm/model_executor/models/qwen2.py", "vllm/model_executor/models/qwen3.py", "TriangleMixBackend",
                 "trianglemix_config", "num_layers"):
        assert must in patch
