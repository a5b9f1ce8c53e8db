"""The drop-in `vectorAdd` executable and the headless full validation pass on a real GPU
(BASELINE.json configs[0] and configs[4])."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "gpu-operator_b200", "validator", "bin")
sys.path.insert(0, os.path.join(ROOT, "tools"))


def _vectoradd(env=None):
    e = {"PATH": os.environ["PATH"]}
    e.update(env or {})
    return subprocess.run([os.path.join(BIN, "vectorAdd")], env=e, capture_output=True, text=True, timeout=300)


def test_vectoradd_binary_keeps_the_sample_contract():
    """No arguments, exit 0 on success (validator/cuda-workload-validation.yaml:19-20; main.go:1049)."""
    r = _vectoradd()
    assert r.returncode == 0, r.stderr
    assert "[Vector addition of 50000 elements]" in r.stdout and "Test PASSED" in r.stdout


def test_vectoradd_binary_bandwidth_run_writes_results(tmp_path):
    r = _vectoradd({"OUTPUT_DIR": str(tmp_path), "NODE_NAME": "node-x", "GPUPROBE_VADD_ELEMS": str(1 << 28)})
    assert r.returncode == 0, r.stderr
    res = json.loads((tmp_path / "gpuprobe-results.json").read_text())
    assert res["pass"] is True and res["node"] == "node-x" and res["vectoradd_check"]["bit_exact"] is True
    g0 = res["vectoradd"][0]
    assert g0["n"] == 1 << 28 and g0["mismatches"] == 0 and g0["gbps"] > 3000
    from oracle import oracle as oc
    assert g0["sum64"] == oc.vectoradd_digest(1 << 28, 0, 0)[0]
    prom = (tmp_path / "gpuprobe-results.prom").read_text()
    assert 'gpu_operator_node_cuda_probe_pass{node="@NODE@"} 1' in prom
    assert 'gpu_operator_node_cuda_hbm_gbps{node="@NODE@",gpu="0"}' in prom


def test_vectoradd_binary_fails_below_the_requested_fraction(tmp_path):
    """GPUPROBE_MIN_HBM_FRAC turns the bandwidth report into a pass/fail criterion (exit 1 -> pod restarts)."""
    r = _vectoradd({"OUTPUT_DIR": str(tmp_path), "GPUPROBE_VADD_ELEMS": str(1 << 28), "GPUPROBE_MIN_HBM_FRAC": "2.0"})
    assert r.returncode == 1
    assert json.loads((tmp_path / "gpuprobe-results.json").read_text())["pass"] is False


def test_headless_full_pass_writes_all_ready_files(tmp_path):
    """driver -> toolkit -> cuda -> plugin in the DaemonSet's order
    (assets/state-operator-validation/0500_daemonset.yaml:29,76,93,116), ready files as before."""
    import full_pass

    res = full_pass.full_pass(vadd_elems=1 << 26, poll_ms=20, out_dir=str(tmp_path))
    assert [s["rc"] for s in res["steps"]] == [0, 0, 0, 0], res["steps"]
    assert {"cuda-ready", "plugin-ready", "toolkit-ready"} <= set(res["ready_files"])
    assert "host-driver-ready" in res["ready_files"] or "driver-ready" in res["ready_files"]
    assert res["results"]["pass"] is True
    assert res["time_to_ready_s"] < 120
