"""The patches under integration/patches/ must apply cleanly to the reference tree they were made against.
Runs only where /root/reference is mounted (the build container); the GPU box skips it."""
import glob
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


PATCHED_FILES = ["validator/main.go", "validator/metrics.go", "validator/Dockerfile",
                 "validator/cuda-workload-validation.yaml", "validator/plugin-workload-validation.yaml",
                 "assets/state-operator-validation/0500_daemonset.yaml"]


@pytest.fixture(scope="module")
def patched_tree(tmp_path_factory):
    """The reference files the patches touch, copied and patched."""
    if not os.path.isdir(REF):
        pytest.skip("/root/reference is only mounted in the build container")
    tmp = tmp_path_factory.mktemp("patched")
    patches = sorted(glob.glob(os.path.join(ROOT, "integration", "patches", "*.patch")))
    assert len(patches) >= 5
    for rel in PATCHED_FILES:
        dst = tmp / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy(os.path.join(REF, rel), dst)
    for p in patches:
        r = subprocess.run(["patch", "-p1", "--forward", "-i", p], cwd=tmp, capture_output=True, text=True)
        assert r.returncode == 0, (p, r.stdout, r.stderr)
    return tmp


def test_patches_apply_to_the_reference(patched_tree):
    tmp_path = patched_tree
    main_go = (tmp_path / "validator" / "main.go").read_text()
    assert main_go.count("forwardProbeEnv(pod)") == 2 and 'case "cuda-workload":' in main_go
    assert "go nm.watchProbeResults()" in (tmp_path / "validator" / "metrics.go").read_text()
    assert "GPUPROBE_VADD_ELEMS" in (tmp_path / "assets" / "state-operator-validation" / "0500_daemonset.yaml").read_text()
    assert "libgpuprobe.so" in (tmp_path / "validator" / "Dockerfile").read_text()


def test_patched_dockerfile_copies_only_from_stages_already_declared(patched_tree):
    """Docker resolves `COPY --from=X` against the stages declared ABOVE the instruction; a name declared later is
    taken for an image and the build fails. Also: nothing refers to the removed sample-builder stage any more, and
    the final image gets both artefacts."""
    import re

    declared = []
    lines = (patched_tree / "validator" / "Dockerfile").read_text().splitlines()
    for n, line in enumerate(lines, 1):
        m = re.match(r"\s*FROM\s+\S+(?:\s+AS\s+(\S+))?", line, re.I)
        if m:
            declared.append(m.group(1))
            continue
        m = re.search(r"COPY\s+--from=(\S+)", line)
        if m:
            assert m.group(1) in declared[:-1], f"line {n}: COPY --from={m.group(1)} before that stage is declared"
    assert "sample-builder" not in "\n".join(lines)
    final = "\n".join(lines[max(i for i, l in enumerate(lines) if l.startswith("FROM ")):])
    assert "COPY --from=gpuprobe-builder /artifacts/vectorAdd /usr/bin/vectorAdd" in final
    assert "COPY --from=gpuprobe-builder /artifacts/libgpuprobe.so /usr/lib64/libgpuprobe.so" in final
    # the image removes every CUDA user library (reference validator/Dockerfile:40), so NCCL is linked in by default
    assert "ARG GPUPROBE_NCCL_STATIC=1" in "\n".join(lines)


def test_patched_workload_pods_mount_the_directory_output_dir_points_at(patched_tree, tmp_path):
    """forwardProbeEnv sets OUTPUT_DIR=/run/nvidia/validations on the workload pod; the patched manifests must mount
    that hostPath (as the validator DaemonSet does, assets/state-operator-validation/0500_daemonset.yaml:50-51,158-162),
    and the mirror's reader must take the patched files as they are: with volume injection off, the results a fake
    vectorAdd writes to $OUTPUT_DIR have to show up in the validator's output directory."""
    import yaml

    ds = yaml.safe_load(open(os.path.join(REF, "assets/state-operator-validation/0500_daemonset.yaml")))
    ds_vol = [v for v in ds["spec"]["template"]["spec"]["volumes"] if v["name"] == "run-nvidia-validations"][0]
    for f in ("cuda-workload-validation.yaml", "plugin-workload-validation.yaml"):
        pod = yaml.safe_load((patched_tree / "validator" / f).read_text())
        vol = [v for v in pod["spec"]["volumes"] if v["name"] == "run-nvidia-validations"][0]
        assert vol == ds_vol
        mounts = pod["spec"]["initContainers"][0]["volumeMounts"]
        assert {"name": "run-nvidia-validations", "mountPath": "/run/nvidia/validations"} in mounts
    fake = tmp_path / "bin"
    fake.mkdir()
    (fake / "vectorAdd").write_text("#!/bin/sh\necho from-the-pod > $OUTPUT_DIR/gpuprobe-results.json\n")
    (fake / "vectorAdd").chmod(0o755)
    validator = os.path.join(ROOT, "gpu-operator_b200", "validator", "bin", "nvidia-validator")
    for comp, extra in (("cuda", {}), ("plugin", {"WITH_WORKLOAD": "true", "GPUPROBE_NODE_CAPACITY_FILE": str(tmp_path / "cap")})):
        (tmp_path / "cap").write_text("nvidia.com/gpu 1\n")
        out = tmp_path / f"v-{comp}"
        env = {"PATH": os.environ["PATH"], "GPUPROBE_POLL_MS": "5", "GPUPROBE_MANIFEST_DIR": str(patched_tree / "validator"),
               "GPUPROBE_BIN_DIR": str(fake), "GPUPROBE_NO_VOLUME_INJECT": "1", **extra}
        r = subprocess.run([validator, "-c", comp, "-n", "node1", "-ns", "ns", "-o", str(out)], env=env, capture_output=True, text=True, timeout=60)
        assert r.returncode == 0, r.stderr
        assert (out / "gpuprobe-results.json").read_text().strip() == "from-the-pod"
        assert "does not mount" not in r.stderr


def test_daemonset_defaults_cannot_block_a_node_without_nccl(patched_tree):
    """Round-1 advice: with GPUPROBE_AR_COUNT on by default and NCCL absent from the image, vectorAdd used to exit 1 on
    every multi-GPU node and cuda-ready was never written. The drop-in now falls back to the library's own allreduce
    kernel on GPUPROBE_ENCCL (unless GPUPROBE_AR_REQUIRE_NCCL=1), in C++ and in the Go binding alike."""
    ds = (patched_tree / "assets" / "state-operator-validation" / "0500_daemonset.yaml").read_text()
    assert "GPUPROBE_AR_COUNT" in ds and "GPUPROBE_AR_REQUIRE_NCCL" not in ds
    cpp = open(os.path.join(ROOT, "gpu-operator_b200", "validator", "vectoradd_main.cpp")).read()
    go = open(os.path.join(ROOT, "integration", "validator", "gpuprobe.go")).read()
    assert "GPUPROBE_ENCCL" in cpp and "gpuprobe_allreduce_i32_auto" in cpp
    assert "C.GPUPROBE_ENCCL" in go and "C.gpuprobe_allreduce_i32_auto" in go


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
