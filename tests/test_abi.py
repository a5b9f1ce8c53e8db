"""The C-ABI boundary on a box without a GPU: the library loads, exports every symbol that
include/gpuprobe.h declares, the ctypes mirror matches the C struct layout, the host-side
generator equals the oracle, and compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import oracle as oc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gpuprobe.h")


def _declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"GPUPROBE_API\s+[\w\s\*]+?\b(gpuprobe_\w+)\s*\(", text)))


def test_header_declares_what_the_binding_binds(gp):
    declared = _declared_symbols()
    assert len(declared) >= 17
    assert sorted(gp.SIGNATURES) == declared


def test_library_exports_every_declared_symbol(gp):
    out = subprocess.run(["nm", "-D", "--defined-only", gp.LIB_PATH], check=True, capture_output=True, text=True).stdout
    exported = set(re.findall(r"\sT\s+(gpuprobe_\w+)", out))
    assert set(_declared_symbols()) <= exported
    # nothing but the C ABI leaks out of the library
    leaked = [l for l in out.splitlines() if " T " in l and "gpuprobe_" not in l]
    assert not leaked, leaked


def test_library_has_no_link_time_cuda_dependency(gp):
    """cudart is static and libcuda/NCCL are dlopen'ed, so the validator's non-GPU components can load it."""
    out = subprocess.run(["readelf", "-d", gp.LIB_PATH], check=True, capture_output=True, text=True).stdout
    needed = re.findall(r"NEEDED.*\[(.*?)\]", out)
    assert not [n for n in needed if re.search(r"cuda|nccl|cublas|torch", n)], needed


def test_library_holds_sm100a_code_only(gp):
    out = subprocess.run(["cuobjdump", "-lelf", gp.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_abi_version(gp):
    m = re.search(r"#define GPUPROBE_ABI_VERSION\s+(\d+)", open(HEADER).read())
    assert gp.lib.gpuprobe_abi_version() == int(m.group(1)) == gp.ABI_VERSION == 2


def test_struct_layout_matches_c(gp, tmp_path):
    """Compile a tiny C program against the header and compare sizeof/offsetof with ctypes."""
    structs = {"gpuprobe_devinfo": gp.DevInfo, "gpuprobe_vadd_cfg": gp.VaddCfg, "gpuprobe_vadd_result": gp.VaddResult,
               "gpuprobe_p2p_cfg": gp.P2PCfg, "gpuprobe_p2p_cell": gp.P2PCell, "gpuprobe_ar_cfg": gp.ArCfg,
               "gpuprobe_ar_result": gp.ArResult}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void){"]
    for cname, ct in structs.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in ct._fields_:
            lines.append(f'printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-o", str(exe), str(src)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, ct in structs.items():
        assert int(got[cname]) == C.sizeof(ct), cname
        for fname, _ in ct._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(ct, fname).offset, f"{cname}.{fname}"


def test_status_codes_match_header(gp):
    text = open(HEADER).read()
    for name in ["OK", "FAIL_VERIFY", "FAIL_BANDWIDTH", "EINVAL", "ENODEV", "EARCH", "ECUDA", "ENOMEM", "ENOPEER", "ENCCL"]:
        m = re.search(rf"#define GPUPROBE_{name}\s+\(?(-?\d+)\)?", text)
        assert m and int(m.group(1)) == getattr(gp, name), name
        assert gp.strerror(getattr(gp, name)) != "unknown gpuprobe status"
    for vid, vname in gp.VARIANT_NAMES.items():
        cname = {"auto": "AUTO", "tma_16k": "TMA"}.get(vname, vname.upper())
        m = re.search(rf"#define GPUPROBE_VADD_{cname}\s+(\d+)", text)
        assert m and int(m.group(1)) == vid, vname


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_host_generator_equals_oracle(gp, kind):
    """gpuprobe_fill_inputs (product host code) against the independent oracle."""
    for first, n, seed in [(0, 50000, 0), (7, 4099, 1), ((1 << 32) - 100, 300, 2), ((1 << 36) + 5, 1001, 3)]:
        a = np.empty(n, dtype=np.float32)
        b = np.empty(n, dtype=np.float32)
        gp.fill_inputs(a, b, n, seed=seed, gen_kind=kind, first=first)
        oa, ob = oc.fill_inputs(n, seed, kind, first)
        assert oc.count_mismatch(a, oa) == 0 and oc.count_mismatch(b, ob) == 0


def test_host_generator_golden(gp, golden):
    kinds = {"unit": 0, "edge": 1, "bits": 2}
    for g in golden["vectoradd"]:
        if g["n"] > 50000:
            continue
        n = g["n"]
        a = np.empty(n, dtype=np.float32)
        b = np.empty(n, dtype=np.float32)
        gp.fill_inputs(a, b, n, seed=g["seed"], gen_kind=kinds[g["kind"]])
        assert a[:4].view(np.uint32).tolist() == g["a_head"][:n]
        assert b[:4].view(np.uint32).tolist() == g["b_head"][:n]


def test_p2p_expected_equals_oracle_and_golden(gp, golden):
    for g in golden["p2p"]:
        assert gp.p2p_expected(g["nwords"], g["seed"], g["src"]) == (g["sum64"], g["xor32"])
    assert gp.p2p_expected(3_000_001, 5, 6, threads=3) == oc.p2p_checksum(3_000_001, 5, 6)
    assert gp.p2p_expected(0, 0, 0) == (0, 0)


def _no_gpu(gp):
    try:
        return gp.device_count() == 0
    except gp.GpuProbeError:
        return True


def test_compute_entry_points_fail_loudly_without_a_gpu(gp):
    """No CPU fallback: on a box without a device every compute entry point returns ENODEV."""
    if not _no_gpu(gp):
        pytest.skip("a GPU is present")
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd(50000)
    assert e.value.code == gp.ENODEV
    a = np.zeros(16, dtype=np.float32)
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd_buffers(a, a, a.copy(), 16)
    assert e.value.code == gp.ENODEV
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd_device(0, 0, 0, 0)
    assert e.value.code == gp.ENODEV
    with pytest.raises(gp.GpuProbeError) as e:
        gp.digest_device(0, 0)
    assert e.value.code == gp.ENODEV
    rc, _ = gp.p2p_matrix(2, 1 << 20, raise_on_fail=False)
    assert rc == gp.ENODEV


def test_bad_arguments_are_rejected_before_touching_a_device(gp):
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd(16, iters=0)
    assert e.value.code == gp.EINVAL
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd(16, variant=99)
    assert e.value.code == gp.EINVAL
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd(16, gen_kind=7)
    assert e.value.code == gp.EINVAL
    rc, _ = gp.p2p_matrix(1, 1 << 20, raise_on_fail=False)
    assert rc == gp.EINVAL
    rc, _ = gp.p2p_matrix(2, 1000, raise_on_fail=False)  # not a multiple of 16
    assert rc == gp.EINVAL
    rc, _ = gp.p2p_matrix(2, 1 << 20, mode=6, raise_on_fail=False)  # modes are 0..5
    assert rc == gp.EINVAL
    with pytest.raises(gp.GpuProbeError) as e:
        gp.allreduce_i32(1, 1024)
    assert e.value.code == gp.EINVAL
    with pytest.raises(gp.GpuProbeError) as e:
        gp.allreduce_i32(2, 0)
    assert e.value.code == gp.EINVAL
