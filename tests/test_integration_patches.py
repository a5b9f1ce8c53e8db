"""The patches under integration/patches/ must apply cleanly to the reference tree they were made against.
Runs only where /root/reference is mounted (the build container); the GPU box skips it."""
import glob
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(REF), reason="/root/reference is only mounted in the build container")
def test_patches_apply_to_the_reference(tmp_path):
    patches = sorted(glob.glob(os.path.join(ROOT, "integration", "patches", "*.patch")))
    assert len(patches) >= 4
    # copy only the files the patches touch
    for rel in ["validator/main.go", "validator/metrics.go", "validator/Dockerfile",
                "assets/state-operator-validation/0500_daemonset.yaml"]:
        dst = tmp_path / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy(os.path.join(REF, rel), dst)
    for p in patches:
        r = subprocess.run(["patch", "-p1", "--forward", "-i", p], cwd=tmp_path, capture_output=True, text=True)
        assert r.returncode == 0, (p, r.stdout, r.stderr)
    main_go = (tmp_path / "validator" / "main.go").read_text()
    assert main_go.count("forwardProbeEnv(pod)") == 2 and 'case "cuda-workload":' in main_go
    assert "go nm.watchProbeResults()" in (tmp_path / "validator" / "metrics.go").read_text()
    assert "GPUPROBE_VADD_ELEMS" in (tmp_path / "assets" / "state-operator-validation" / "0500_daemonset.yaml").read_text()
    assert "libgpuprobe.so" in (tmp_path / "validator" / "Dockerfile").read_text()


def test_go_sources_reference_only_declared_c_symbols():
    """Every C.gpuprobe_* call in the (uncompiled) Go binding names a symbol that include/gpuprobe.h declares."""
    import re

    header = open(os.path.join(ROOT, "include", "gpuprobe.h")).read()
    declared = set(re.findall(r"GPUPROBE_API\s+[\w\s\*]+?\b(gpuprobe_\w+)\s*\(", header))
    structs = set(re.findall(r"typedef struct (gpuprobe_\w+)", header))
    consts = set(re.findall(r"#define (GPUPROBE_\w+)", header))
    go = open(os.path.join(ROOT, "integration", "validator", "gpuprobe.go")).read()
    used_fn = set(re.findall(r"C\.(gpuprobe_\w+)\(", go))
    used_ty = set(re.findall(r"C\.(gpuprobe_\w+)\{", go)) | set(re.findall(r"(?:var \w+ |\[\])C\.(gpuprobe_\w+)", go))
    used_const = set(re.findall(r"C\.(GPUPROBE_\w+)", go))
    assert used_fn and used_fn <= declared, used_fn - declared
    assert used_ty <= structs, used_ty - structs
    assert used_const <= consts, used_const - consts
