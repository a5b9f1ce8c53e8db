"""The host-side mirror of the reference validator (gpu-operator_b200/validator): same flags,
env fallbacks, error text, status files and retry behaviour as validator/main.go for the
`cuda` and `plugin` components, with the Kubernetes hops replaced by a child process.
CPU only: the workload is a shell command, no GPU code runs here.
(The reference has no tests for validator/ at all -- SURVEY.md section 4 -- so these are written
against the behaviour of validator/main.go, cited per test.)"""
import os
import signal
import socket
import subprocess
import time
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "gpu-operator_b200", "validator", "bin")
VALIDATOR = os.path.join(BIN, "nvidia-validator")

CUDA_MANIFEST = """kind: Pod
metadata:
  labels:
    app: nvidia-cuda-validator
  generateName: nvidia-cuda-validator-
spec:
  restartPolicy: {restart}
  initContainers:
  - name: cuda-validation
    command: ['sh', '-c']
    args: ["{cmd}"]
    env:
    - name: NVIDIA_VISIBLE_DEVICES
      value: "all"
"""
PLUGIN_MANIFEST = """kind: Pod
metadata:
  labels:
    app: nvidia-device-plugin-validator
  generateName: nvidia-device-plugin-validator-
spec:
  restartPolicy: OnFailure
  initContainers:
  - name: plugin-validation
    command: ['sh', '-c']
    args: ["{cmd}"]
    resources:
      limits:
        "FILLED_BY_VALIDATOR": 1
"""


@pytest.fixture(scope="module", autouse=True)
def built():
    if not os.path.exists(VALIDATOR):
        subprocess.run(["make", "-C", os.path.dirname(BIN)], check=True, capture_output=True)


def run(args=(), env=None, timeout=60):
    e = {"PATH": os.environ["PATH"], "GPUPROBE_POLL_MS": "5", "GPUPROBE_KEEP_RETRY_COUNTS": "1", "SLEEP_INTERVAL_SECONDS": "0"}
    e.update(env or {})
    return subprocess.run([VALIDATOR, *args], env=e, capture_output=True, text=True, timeout=timeout)


def manifests(tmp_path, cuda_cmd="true", plugin_cmd="true", restart="OnFailure"):
    d = tmp_path / "manifests"
    d.mkdir(exist_ok=True)
    (d / "cuda-workload-validation.yaml").write_text(CUDA_MANIFEST.format(cmd=cuda_cmd, restart=restart))
    (d / "plugin-workload-validation.yaml").write_text(PLUGIN_MANIFEST.format(cmd=plugin_cmd))
    return str(d)


# ---- validateFlags (validator/main.go:339-370): same messages, exit status 1, "Error: " prefix (:321-326)
@pytest.mark.parametrize("args,env,msg", [
    ([], {}, "invalid -c <component-name> flag: must not be empty string"),
    (["-c", "bogus"], {}, "invalid -c <component-name> flag value: bogus"),
    (["--component=plugin"], {}, "invalid -n <node-name> flag: must not be empty string for plugin validation"),
    (["-c", "plugin", "-n", "node1"], {}, "invalid -ns <namespace> flag: must not be empty string for plugin validation"),
    (["-c", "cuda"], {}, "invalid -ns <namespace> flag: must not be empty string for cuda validation"),
    (["-c", "metrics", "-n", "node1"], {}, "invalid -p <port> flag: must not be empty or 0 for the metrics component"),
    (["-c", "metrics", "-p", "8000"], {}, "invalid -n <node-name> flag: must not be empty string for metrics exporter"),
    (["-c", "vgpu-devices"], {}, "invalid -n <node-name> flag: must not be empty string for vgpu-devices validation"),
    ([], {"COMPONENT": "cuda"}, "invalid -ns <namespace> flag: must not be empty string for cuda validation"),
    (["--no-such-flag"], {}, "flag provided but not defined: -no-such-flag"),
])
def test_flag_validation_messages(args, env, msg):
    r = run(args, env)
    assert r.returncode == 1
    assert r.stderr.strip().splitlines()[-1] == "Error: " + msg


def test_cuda_component_creates_ready_file_and_removes_a_stale_one_first(tmp_path):
    """CUDA.validate (main.go:1176-1211): delete cuda-ready, run the workload, create cuda-ready."""
    out = tmp_path / "validations"
    out.mkdir()
    (out / "cuda-ready").write_text("stale")
    probe = tmp_path / "seen"
    # the workload records whether the stale file was already gone when it ran
    cmd = f"test ! -e {out}/cuda-ready && touch {probe}"
    r = run(["-c", "cuda", "--namespace", "gpu-operator", "-o", str(out)],
            {"GPUPROBE_MANIFEST_DIR": manifests(tmp_path, cuda_cmd=cmd), "NODE_NAME": "node1"})
    assert r.returncode == 0, r.stderr
    assert probe.exists()
    assert (out / "cuda-ready").exists() and (out / "cuda-ready").read_text() == ""  # os.Create: empty file
    assert "have run successfully" in r.stderr  # main.go:1055


def test_env_fallbacks_and_probe_env_forwarding(tmp_path):
    """All configuration may come from env (0500_daemonset.yaml:94-109 runs `sh -c nvidia-validator`);
    GPUPROBE_* knobs are forwarded into the workload (new; main.go:1225-1245 forwards only image settings)."""
    out = tmp_path / "v"
    dump = tmp_path / "env.txt"
    r = run([], {"COMPONENT": "cuda", "OPERATOR_NAMESPACE": "ns", "OUTPUT_DIR": str(out), "NODE_NAME": "n1",
                 "GPUPROBE_VADD_ELEMS": "1048576", "VALIDATOR_IMAGE": "img:1",
                 "GPUPROBE_MANIFEST_DIR": manifests(tmp_path, cuda_cmd=f"env > {dump}")})
    assert r.returncode == 0, r.stderr
    env = dict(l.split("=", 1) for l in dump.read_text().splitlines() if "=" in l)
    assert env["GPUPROBE_VADD_ELEMS"] == "1048576"
    assert env["NVIDIA_VISIBLE_DEVICES"] == "all"  # from the pod spec (cuda-workload-validation.yaml:21-23)
    assert env["OUTPUT_DIR"] == str(out) and env["NODE_NAME"] == "n1"
    assert (out / "cuda-ready").exists()  # start() created the output dir (main.go:451-455)


def test_failing_workload_is_restarted_then_succeeds(tmp_path):
    """restartPolicy: OnFailure (cuda-workload-validation.yaml:13): a non-zero exit restarts the container."""
    out = tmp_path / "v"
    counter = tmp_path / "count"
    cmd = f"echo x >> {counter}; test $(wc -l < {counter}) -ge 3"
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out)], {"GPUPROBE_MANIFEST_DIR": manifests(tmp_path, cuda_cmd=cmd)})
    assert r.returncode == 0, r.stderr
    assert counter.read_text().count("x") == 3
    assert (out / "cuda-ready").exists()


def test_workload_that_never_succeeds_gives_up_like_waitForPod(tmp_path):
    """waitForPod (main.go:1042-1059): 60 polls, then "gave up waiting for pod ... to be available";
    the error is wrapped by start() (main.go:484) and no ready file is written."""
    out = tmp_path / "v"
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out)], {"GPUPROBE_MANIFEST_DIR": manifests(tmp_path, cuda_cmd="false")})
    assert r.returncode == 1
    last = r.stderr.strip().splitlines()[-1]
    assert last.startswith("Error: error validating cuda workload: gave up waiting for pod nvidia-cuda-validator-")
    assert last.endswith(" to be available")
    assert not (out / "cuda-ready").exists()
    assert r.stderr.count("is curently in Pending phase") == 60  # sic (main.go:1051), podCreationWaitRetries


def test_manifest_with_block_style_lists_and_comments(tmp_path):
    """loadPodSpec (main.go:1061-1081) decodes any YAML; the mirror's reader must at least take both list styles."""
    d = tmp_path / "manifests"
    d.mkdir()
    marker = tmp_path / "ran"
    (d / "cuda-workload-validation.yaml").write_text(f"""apiVersion: v1
kind: Pod          # a comment
metadata:
    labels:
        app: nvidia-cuda-validator
    generateName: "nvidia-cuda-validator-"
spec:
    restartPolicy: OnFailure
    initContainers:
    - name: cuda-validation
      image: 'x'
      command:
      - sh
      - "-c"
      args:
      - "echo $PROBE_MARK > {marker}"
      env:
      - name: PROBE_MARK
        value: "block-style"
    containers:
    - name: nvidia-cuda-validator
      command: ['sh', '-c']
      args: ["echo done"]
""")
    out = tmp_path / "v"
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out)], {"GPUPROBE_MANIFEST_DIR": str(d)})
    assert r.returncode == 0, r.stderr
    assert marker.read_text().strip() == "block-style"
    assert (out / "cuda-ready").exists()


def test_manifest_comment_marker_inside_quotes_is_data(tmp_path):
    """A '#' inside a quoted string is part of the string, not the start of a comment."""
    marker = tmp_path / "ran"
    d = manifests(tmp_path)
    text = CUDA_MANIFEST.format(cmd=f"echo 'a #b' > {marker}", restart="OnFailure")
    (tmp_path / "manifests" / "cuda-workload-validation.yaml").write_text(
        text.replace("restartPolicy: OnFailure", "restartPolicy: OnFailure   # a real comment"))
    out = tmp_path / "v"
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out)], {"GPUPROBE_MANIFEST_DIR": d})
    assert r.returncode == 0, r.stderr
    assert marker.read_text().strip() == "a #b"


@pytest.mark.parametrize("snippet,what", [
    ('    args: |\n      vectorAdd\n', "a block scalar"),
    ('    args: >-\n      vectorAdd\n', "a block scalar"),
    ('    args: *defaultargs\n', "an alias"),
    ('    args: &a ["true"]\n', "an anchor"),
    ('    args: ["true"]\n    securityContext: {privileged: true}\n', "a flow mapping"),
    ('    args: ["true"]\n    <<: *base\n', "a merge key"),
    ('    args: ["true"]\n---\nkind: Pod\n', "a second YAML document"),
    ('    args: ["true"]\n\tenv: []\n', "tab indentation"),
])
def test_manifest_constructs_the_reader_cannot_take_are_named(tmp_path, snippet, what):
    """The mirror's loadPodSpec is a reader for the YAML subset of the two shipped manifests (the reference decodes
    any YAML, main.go:1061-1081); what it cannot read it refuses with the construct and the line, never silently."""
    d = tmp_path / "manifests"
    d.mkdir()
    (d / "cuda-workload-validation.yaml").write_text(
        "kind: Pod\nmetadata:\n  generateName: x-\nspec:\n  restartPolicy: Never\n  initContainers:\n  - name: cuda-validation\n" + snippet)
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(tmp_path / "v")], {"GPUPROBE_MANIFEST_DIR": str(d)})
    assert r.returncode == 1
    last = r.stderr.strip().splitlines()[-1]
    assert what in last and "is not supported by the validator's manifest reader" in last
    assert "cuda-workload-validation.yaml:" in last  # path:line


VOLUMES = """    volumeMounts:
      - name: run-nvidia-validations
        mountPath: /run/nvidia/validations
  volumes:
    - name: run-nvidia-validations
      hostPath:
        path: /run/nvidia/validations
        type: DirectoryOrCreate
"""


@pytest.mark.parametrize("mounted,inject", [(True, False), (False, True), (False, False)])
def test_probe_results_leave_the_pod_only_through_the_validations_mount(tmp_path, mounted, inject):
    """The results file is written inside the workload pod; the metrics exporter reads it from the node's
    /run/nvidia/validations. It gets there only if the pod mounts that hostPath (the reference's workload pods declare
    no volumes at all, validator/cuda-workload-validation.yaml:1-35). Three cases: the manifest mounts it; it does not
    and the validator adds the mount (what forwardProbeEnv does in the Go patch); neither -- validation still passes,
    but loudly without results."""
    d = tmp_path / "manifests"
    d.mkdir()
    text = CUDA_MANIFEST.format(cmd="echo pass-true > $OUTPUT_DIR/gpuprobe-results.json || echo cannot write results >&2", restart="Never")
    (d / "cuda-workload-validation.yaml").write_text(text + (VOLUMES if mounted else ""))
    out = tmp_path / "v"
    env = {"GPUPROBE_MANIFEST_DIR": str(d)}
    if not inject:
        env["GPUPROBE_NO_VOLUME_INJECT"] = "1"
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out)], env)
    assert r.returncode == 0, r.stderr
    assert (out / "cuda-ready").exists()
    if mounted or inject:
        assert (out / "gpuprobe-results.json").read_text().strip() == "pass-true"
        assert "does not mount" not in r.stderr
    else:
        assert not (out / "gpuprobe-results.json").exists()
        assert "cuda workload pod does not mount /run/nvidia/validations" in r.stderr


def test_shipped_headless_manifests_mount_the_validations_dir():
    for f in ("cuda-workload-validation.yaml", "plugin-workload-validation.yaml"):
        text = open(os.path.join(ROOT, "gpu-operator_b200", "validator", "manifests", f)).read()
        assert "mountPath: /run/nvidia/validations" in text and "path: /run/nvidia/validations" in text


def test_missing_manifest_is_an_error(tmp_path):
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(tmp_path / "v")], {"GPUPROBE_MANIFEST_DIR": str(tmp_path / "nowhere")})
    assert r.returncode == 1 and "error validating cuda workload: open " in r.stderr


def test_plugin_without_workload_needs_only_discovered_resources(tmp_path):
    """Plugin.validate (main.go:803-845) with WITH_WORKLOAD unset (the chart's default, values.yaml:54-57)."""
    out = tmp_path / "v"
    cap = tmp_path / "capacity"
    cap.write_text("cpu 64\nnvidia.com/gpu 8\n")
    r = run(["-c", "plugin", "-n", "node1", "-ns", "ns", "-o", str(out)], {"GPUPROBE_NODE_CAPACITY_FILE": str(cap)})
    assert r.returncode == 0, r.stderr
    assert (out / "plugin-ready").exists()


def test_plugin_gives_up_when_no_gpu_resource_appears(tmp_path):
    """validateGPUResource (main.go:1102-1122): 30 retries, then this exact error."""
    out = tmp_path / "v"
    cap = tmp_path / "capacity"
    cap.write_text("cpu 64\nnvidia.com/gpu 0\n")
    r = run(["-c", "plugin", "-n", "node1", "-ns", "ns", "-o", str(out)], {"GPUPROBE_NODE_CAPACITY_FILE": str(cap)})
    assert r.returncode == 1
    assert r.stderr.strip().splitlines()[-1] == "Error: error validating plugin installation: GPU resources are not discovered by the node"
    assert r.stderr.count("GPU resources are not yet discovered by the node, retry:") == 30
    assert not (out / "plugin-ready").exists()


def test_plugin_workload_prefers_mig_resource_and_sees_one_gpu(tmp_path):
    """getGPUResourceName (main.go:1144-1161): a MIG resource wins over nvidia.com/gpu; the workload
    pod requests exactly one of it (main.go:973-983), so the child sees a single device."""
    out = tmp_path / "v"
    cap = tmp_path / "capacity"
    cap.write_text("nvidia.com/gpu 4\nnvidia.com/mig-1g.23gb 7\n")
    dump = tmp_path / "env.txt"
    r = run(["-c", "plugin", "-n", "node1", "-ns", "ns", "-o", str(out), "--with-workload"],
            {"GPUPROBE_NODE_CAPACITY_FILE": str(cap), "GPUPROBE_MANIFEST_DIR": manifests(tmp_path, plugin_cmd=f"env > {dump}")})
    assert r.returncode == 0, r.stderr
    env = dict(l.split("=", 1) for l in dump.read_text().splitlines() if "=" in l)
    assert env["CUDA_VISIBLE_DEVICES"] == "0"
    assert (out / "plugin-ready").exists()
    assert "nvidia-device-plugin-validator-" in r.stderr


def test_cleanup_all_recreates_the_output_dir(tmp_path):
    """start() (main.go:439-455): --cleanup-all removes the whole output dir, then it is created again."""
    out = tmp_path / "v"
    out.mkdir()
    (out / "driver-ready").write_text("")
    (out / "junk").write_text("x")
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out), "-r"], {"GPUPROBE_MANIFEST_DIR": manifests(tmp_path)})
    assert r.returncode == 0, r.stderr
    assert sorted(p.name for p in out.iterdir()) == ["cuda-ready"]


def test_components_outside_the_path_are_named_but_not_built(tmp_path):
    r = run(["-c", "mofed", "-o", str(tmp_path / "v")])
    assert r.returncode == 1 and "not part of the CUDA-workload path" in r.stderr


def test_sigterm_exits_1_like_handleSignal(tmp_path):
    """handleSignal (main.go:329-337): log.Fatalf -> exit status 1."""
    pidfile = tmp_path / "workload.pid"
    e = {"PATH": os.environ["PATH"], "GPUPROBE_MANIFEST_DIR": manifests(tmp_path, cuda_cmd=f"echo $$ > {pidfile}; exec sleep 30")}
    errlog = tmp_path / "stderr.log"
    with open(errlog, "w") as ef:
        p = subprocess.Popen([VALIDATOR, "-c", "cuda", "-ns", "ns", "-o", str(tmp_path / "v")], env=e, stderr=ef)
        for _ in range(100):
            if pidfile.exists() and pidfile.read_text().strip():
                break
            time.sleep(0.05)
        p.send_signal(signal.SIGTERM)
        p.wait(timeout=10)
    try:
        os.kill(int(pidfile.read_text()), signal.SIGKILL)  # the workload outlives the validator, as a pod would
    except (ProcessLookupError, ValueError):
        pass
    assert p.returncode == 1
    assert "Exiting due to signal [terminated] notification for pid" in errlog.read_text()


def test_reference_manifests_parse(tmp_path):
    """The same reader handles the reference's own workload manifests (validator/*-workload-validation.yaml)."""
    ref = "/root/reference/validator"
    if not os.path.isdir(ref):
        pytest.skip("/root/reference is only mounted in the build container")
    out = tmp_path / "v"
    dump = tmp_path / "which"
    fake = tmp_path / "bin"
    fake.mkdir()
    (fake / "vectorAdd").write_text(f"#!/bin/sh\necho ran > {dump}\n")
    (fake / "vectorAdd").chmod(0o755)
    r = run(["-c", "cuda", "-ns", "ns", "-o", str(out)], {"GPUPROBE_MANIFEST_DIR": ref, "GPUPROBE_BIN_DIR": str(fake)})
    assert r.returncode == 0, r.stderr
    assert dump.read_text().strip() == "ran" and (out / "cuda-ready").exists()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_metrics_component_serves_reference_gauges_plus_probe_gauges(tmp_path):
    """--component=metrics (validator/metrics.go:73-157, 304-320): the *_ready gauges follow the status
    files; the probe's achieved-GB/s gauges come from gpuprobe-results.prom in the same directory."""
    out = tmp_path / "v"
    out.mkdir()
    port = _free_port()
    # the watchers run on the reference's cadence (status files and probe results every 30 s, driver and PCI every 60 s,
    # metrics.go:40-47); the test shrinks a "second" of those delays to 1 ms
    e = {"PATH": os.environ["PATH"], "GPUPROBE_WATCH_MS_PER_SECOND": "1"}
    errlog = open(tmp_path / "metrics.stderr", "w")
    p = subprocess.Popen([VALIDATOR, "-c", "metrics", "-n", "node-a", "-p", str(port), "-o", str(out)], env=e,
                         stderr=errlog, text=True)
    try:
        def scrape():
            for _ in range(100):
                try:
                    return urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics", timeout=2).read().decode()
                except OSError:
                    time.sleep(0.05)
            raise AssertionError("metrics endpoint did not come up")

        m = scrape()
        assert 'gpu_operator_node_cuda_ready{node="node-a"} 0' in m
        assert 'gpu_operator_node_plugin_ready{node="node-a"} 0' in m
        assert "# TYPE gpu_operator_node_driver_ready gauge" in m
        assert "gpu_operator_node_metrics_ready_ts_seconds" in m
        # all ten gauges of validator/metrics.go:77-155 are exported, each labelled {node}
        for name in ["metrics_ready_ts_seconds", "driver_ready", "toolkit_ready", "plugin_ready", "cuda_ready",
                     "driver_validation", "driver_validation_last_success_ts_seconds", "device_plugin_devices_total",
                     "device_plugin_validation_last_success_ts_seconds"]:
            assert f'gpu_operator_node_{name}{{node="node-a"}}' in m, name
        assert 'gpu_operator_nvidia_pci_devices_total{node="node-a"}' in m
        (out / "cuda-ready").write_text("")
        (out / "host-driver-ready").write_text("")  # metrics.go:169-173: either driver file opens the barrier
        (out / "gpuprobe-results.prom").write_text(
            '# TYPE gpu_operator_node_cuda_hbm_gbps gauge\ngpu_operator_node_cuda_hbm_gbps{node="@NODE@",gpu="0"} 7126.5\n')
        for _ in range(200):  # the gauges follow at the watchers' next pass, not at the scrape
            m = scrape()
            if 'gpu_operator_node_cuda_hbm_gbps{node="node-a",gpu="0"} 7126.5' in m and 'cuda_ready{node="node-a"} 1' in m \
                    and 'driver_ready{node="node-a"} 1' in m:
                break
            time.sleep(0.02)
        assert 'gpu_operator_node_cuda_ready{node="node-a"} 1' in m
        assert 'gpu_operator_node_driver_ready{node="node-a"} 1' in m
        assert 'gpu_operator_node_cuda_hbm_gbps{node="node-a",gpu="0"} 7126.5' in m
        (out / "cuda-ready").unlink()
        for _ in range(200):
            m = scrape()
            if 'cuda_ready{node="node-a"} 0' in m:
                break
            time.sleep(0.02)
        assert 'gpu_operator_node_cuda_ready{node="node-a"} 0' in m
        with pytest.raises(urllib.error.HTTPError):
            urllib.request.urlopen(f"http://127.0.0.1:{port}/nope", timeout=2)
    finally:
        p.terminate()
        p.wait(timeout=5)
        errlog.close()
    log = (tmp_path / "metrics.stderr").read_text()
    # transitions are logged the way watchStatusFile logs them (metrics.go:176-186)
    assert "metrics: StatusFile: watching cuda-ready" in log
    assert "metrics: StatusFile: 'cuda-ready' is ready" in log and "metrics: StatusFile: 'cuda-ready' is not ready" in log
    assert "Running the metrics server, listening on :%d/metrics" % port in log


def test_driver_component_follows_the_reference_outside_a_driver_container(tmp_path):
    """Driver.runValidation (main.go:609-628) on a box with neither /host nor /run/nvidia/driver: the reference asks
    for /run/nvidia/validations/.driver-ctr-ready first and fails there; so does the mirror."""
    if os.path.exists("/host/usr/bin/nvidia-smi") or os.path.exists("/run/nvidia/validations/.driver-ctr-ready"):
        pytest.skip("this box looks like a driver container")
    out = tmp_path / "v"
    r = run(["-c", "driver", "-o", str(out)])
    assert r.returncode == 1
    assert r.stderr.strip().splitlines()[-1].startswith(
        "Error: error validating driver installation: error checking driver container status: exit status")
    assert not (out / "driver-ready").exists() and not (out / "host-driver-ready").exists()


def test_driver_component_headless_switch_checks_the_local_driver(tmp_path):
    """GPUPROBE_HEADLESS=1 (tools/full_pass.py): a bare `nvidia-smi` stands in and the result counts as a host driver."""
    if os.path.exists("/host/usr/bin/nvidia-smi") or os.path.exists("/run/nvidia/driver/usr/bin/nvidia-smi"):
        pytest.skip("this box looks like a driver container")
    fake = tmp_path / "bin"
    fake.mkdir()
    (fake / "nvidia-smi").write_text("#!/bin/sh\necho fake driver ok\n")
    (fake / "nvidia-smi").chmod(0o755)
    out = tmp_path / "v"
    r = run(["-c", "driver", "-o", str(out)], {"GPUPROBE_HEADLESS": "1", "PATH": f"{fake}:{os.environ['PATH']}"})
    assert r.returncode == 0, r.stderr
    assert (out / "host-driver-ready").exists() and not (out / "driver-ready").exists()
    (fake / "nvidia-smi").write_text("#!/bin/sh\nexit 9\n")
    r = run(["-c", "driver", "-o", str(out)], {"GPUPROBE_HEADLESS": "1", "PATH": f"{fake}:{os.environ['PATH']}"})
    assert r.returncode == 1 and "error validating driver installation: exit status 9" in r.stderr
    assert not (out / "host-driver-ready").exists()  # deleted at the start of the run (main.go:639-649)
