"""Parity of the CUDA vectorAdd path against the oracle, through the C ABI, on a real B200.

Bit-exact is the bar: fp32 a+b rounds once, so the GPU must reproduce the host loop bit for
bit (BASELINE.json north_star).  Sizes up to a few 10^7 are compared element by element /
digest against the oracle and the committed golden fixtures; the full 2^32 size is covered by
size-independent properties (sampled slices, checksum of checksums, idempotence).
"""
import numpy as np
import pytest

from oracle import oracle as oc

pytestmark = pytest.mark.gpu

KINDS = {"unit": 0, "edge": 1, "bits": 2}


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available(), "these tests need a GPU"
    torch.cuda.set_device(0)
    return torch


@pytest.fixture(scope="module")
def variants(gp):
    return [v for v in range(1, gp.VADD_NUM_VARIANTS)]


def test_device_is_b200(gp):
    info = gp.device_info(0)
    assert info.cc_major == 10, f"sm_{info.cc_major}{info.cc_minor}"
    assert info.sm_count >= 100


def test_reference_configuration_50000(gp, golden):
    """BASELINE.json configs[0]: 50,000 fp32 elements, the size the sample binary uses."""
    res = gp.vectoradd(50_000, iters=20, warmup=3, seed=0, gen_kind=gp.GEN_UNIT, verify=3)
    g = next(x for x in golden["vectoradd"] if x["kind"] == "unit" and x["seed"] == 0 and x["n"] == 50_000)
    assert (res.sum64, res.xor32) == (g["sum64"], g["xor32"])
    assert res.mismatches == 0 and res.n == 50_000 and res.algo_bytes == 600_000
    assert res.l2_resident == 1


def test_golden_digests_every_variant(gp, golden, variants):
    for g in golden["vectoradd"]:
        for v in variants:
            res = gp.vectoradd(g["n"], iters=1, warmup=0, seed=g["seed"], gen_kind=KINDS[g["kind"]], variant=v, verify=3)
            assert (res.sum64, res.xor32) == (g["sum64"], g["xor32"]), (g["kind"], g["seed"], g["n"], gp.VARIANT_NAMES[v])
            assert res.mismatches == 0


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_elementwise_against_host_loop_every_variant(gp, torch_cuda, variants, kind):
    """Host inputs from the oracle generator -> device -> every kernel variant -> compare all of C."""
    torch = torch_cuda
    # sizes straddle every tile size (2048 .. 8192 elements) and leave ragged tails
    for n in [1, 5, 2047, 2048, 2049, 4096 * 3 + 17, 8192 * 5, 1_000_003]:
        a, b = oc.fill_inputs(n, seed=11, kind=kind)
        want = oc.vectoradd(a, b)
        da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        for v in variants:
            dc = torch.full((n,), float("nan"), device="cuda", dtype=torch.float32)
            gp.vectoradd_device(da, db, dc, n, variant=v, stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            got = dc.cpu().numpy()
            bad = oc.count_mismatch(got, want)
            assert bad == 0, (n, gp.VARIANT_NAMES[v], bad)


def test_unaligned_device_pointers_take_the_scalar_path(gp, torch_cuda):
    torch = torch_cuda
    n = 100_001
    a, b = oc.fill_inputs(n + 3, seed=4, kind=2)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    dc = torch.zeros(n + 3, device="cuda", dtype=torch.float32)
    for off_a, off_b, off_c in [(1, 0, 0), (0, 0, 1), (1, 2, 3), (3, 3, 3)]:
        dc.zero_()
        gp.vectoradd_device(da.data_ptr() + 4 * off_a, db.data_ptr() + 4 * off_b, dc.data_ptr() + 4 * off_c, n,
                            stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        got = dc.cpu().numpy()[off_c:off_c + n]
        want = oc.vectoradd(np.ascontiguousarray(a[off_a:off_a + n]), np.ascontiguousarray(b[off_b:off_b + n]))
        assert oc.count_mismatch(np.ascontiguousarray(got), want) == 0
        # nothing outside [off_c, off_c+n) was written
        full = dc.cpu().numpy()
        assert (full[:off_c] == 0).all() and (full[off_c + n:] == 0).all()


def test_no_variant_writes_outside_its_output(gp, torch_cuda, variants):
    """Guard bands around C (compute-sanitizer is closed on this pool, so bounds are checked by sentinel):
    every variant, sizes that end mid-tile, must leave 4 KiB on both sides of C untouched."""
    torch = torch_cuda
    pad = 1024
    s = torch.cuda.current_stream().cuda_stream
    for n in [1, 1023, 1025, 4096 + 7, 8192 * 3 + 1, (1 << 20) - 3]:
        a, b = oc.fill_inputs(n, seed=13, kind=0)
        want = oc.vectoradd(a, b)
        da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        for v in variants:
            buf = torch.full((n + 2 * pad,), -12345.0, device="cuda", dtype=torch.float32)
            gp.vectoradd_device(da, db, buf.data_ptr() + 4 * pad, n, variant=v, stream=s)
            torch.cuda.synchronize()
            host = buf.cpu().numpy()
            assert (host[:pad] == -12345.0).all() and (host[pad + n:] == -12345.0).all(), (n, gp.VARIANT_NAMES[v])
            assert oc.count_mismatch(np.ascontiguousarray(host[pad:pad + n]), want) == 0, (n, gp.VARIANT_NAMES[v])


def test_edge_table_exact(gp, torch_cuda, golden):
    torch = torch_cuda
    t = golden["edge_table"]
    ea = np.ascontiguousarray(np.repeat(np.array(t["a"], dtype=np.uint32), 16).view(np.float32))
    eb = np.ascontiguousarray(np.tile(np.array(t["b"], dtype=np.uint32), 16).view(np.float32))
    # tile the 256 pairs so that the vector kernels (not only the scalar tail) see them
    reps = 64
    a, b = np.tile(ea, reps), np.tile(eb, reps)
    c = np.empty_like(a)
    gp.vectoradd_buffers(a, b, c, a.size)
    assert c.view(np.uint32)[:256].tolist() == t["c_row_major_a_by_b"]
    assert (c.view(np.uint32).reshape(reps, 256) == np.array(t["c_row_major_a_by_b"], dtype=np.uint32)).all()


def test_matches_torch_add_bit_for_bit(gp, torch_cuda):
    """Independent cross-check on the device."""
    torch = torch_cuda
    n = (1 << 22) + 5
    da = torch.empty(n, device="cuda", dtype=torch.float32)
    db = torch.empty_like(da)
    dc = torch.empty_like(da)
    gp.fill_inputs_device(da, db, n, seed=9, gen_kind=gp.GEN_BITS, stream=torch.cuda.current_stream().cuda_stream)
    gp.vectoradd_device(da, db, dc, n, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ref = torch.add(da, db)
    assert torch.equal(dc.view(torch.int32), ref.view(torch.int32))


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_device_generator_equals_oracle(gp, torch_cuda, kind):
    torch = torch_cuda
    for first, n in [(0, 50_000), (3, 4097), ((1 << 32) - 1000, 5001), ((1 << 34) + 1, 333)]:
        da = torch.empty(n, device="cuda", dtype=torch.float32)
        db = torch.empty_like(da)
        gp.fill_inputs_device(da, db, n, seed=2, gen_kind=kind, first=first, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        oa, ob = oc.fill_inputs(n, 2, kind, first)
        assert oc.count_mismatch(da.cpu().numpy(), oa) == 0
        assert oc.count_mismatch(db.cpu().numpy(), ob) == 0
    # unaligned destination
    da = torch.empty(1001, device="cuda", dtype=torch.float32)
    gp.fill_inputs_device(da.data_ptr() + 4, None, 1000, seed=2, gen_kind=kind, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert oc.count_mismatch(np.ascontiguousarray(da.cpu().numpy()[1:]), oc.fill_inputs(1000, 2, kind)[0]) == 0


def test_digest_device_equals_oracle(gp, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(5)
    for n in [0, 1, 3, 255, 1024, 100_003, (1 << 22) + 1]:
        w = rng.integers(0, 2**32, size=n + 1, dtype=np.uint32)
        d = torch.from_numpy(w.view(np.int32)).cuda()
        assert gp.digest_device(d, n) == oc.digest(w[:n].view(np.float32)), n
        if n:  # unaligned start
            assert gp.digest_device(d.data_ptr() + 4, n) == oc.digest(np.ascontiguousarray(w[1:n + 1]).view(np.float32)), n


def test_verify_device_counts_injected_corruption(gp, torch_cuda):
    torch = torch_cuda
    n = 1 << 20
    da = torch.empty(n, device="cuda", dtype=torch.float32)
    db = torch.empty_like(da)
    dc = torch.empty_like(da)
    s = torch.cuda.current_stream().cuda_stream
    gp.fill_inputs_device(da, db, n, seed=3, gen_kind=0, stream=s)
    gp.vectoradd_device(da, db, dc, n, stream=s)
    assert gp.verify_device(dc, n, seed=3, gen_kind=0, stream=s) == 0
    dc.view(torch.int32)[12345] ^= 1  # one flipped mantissa bit
    dc.view(torch.int32)[n - 1] ^= 1 << 31
    assert gp.verify_device(dc, n, seed=3, gen_kind=0, stream=s) == 2


def test_host_buffers_pageable_and_pinned(gp):
    """gpuprobe_vectoradd_buffers: the sample's h_A/h_B -> device -> h_C round trip; the caller's
    host loop (here the oracle) does the compare. Sizes cover empty, one chunk and several chunks."""
    for n in [0, 1, 50_000, (8 << 20) + 12_345, (20 << 20) + 3]:
        a, b = oc.fill_inputs(n, seed=6, kind=2)
        want = oc.vectoradd(a, b, threads=oc.host_threads())
        c = np.full(n, np.nan, dtype=np.float32)
        gp.vectoradd_buffers(a, b, c, n)  # pageable
        assert oc.count_mismatch(c, want) == 0, n
        if n:
            pa, oa_ = gp.pinned_array(n)
            pb, ob_ = gp.pinned_array(n)
            pc, oc_ = gp.pinned_array(n)
            try:
                pa[:] = a
                pb[:] = b
                pc[:] = np.nan
                gp.vectoradd_buffers(pa, pb, pc, n)  # pinned: direct copies
                assert oc.count_mismatch(np.ascontiguousarray(pc), want) == 0, n
            finally:
                del pa, pb, pc
                oa_.free(), ob_.free(), oc_.free()


def test_empty_input_is_ok(gp):
    res = gp.vectoradd(0, iters=1, warmup=0)
    assert res.n == 0 and (res.sum64, res.xor32) == (0, 0) and res.mismatches == 0


def test_idempotent(gp):
    r1 = gp.vectoradd(1 << 24, iters=3, warmup=1, seed=8, gen_kind=2)
    r2 = gp.vectoradd(1 << 24, iters=1, warmup=0, seed=8, gen_kind=2)
    assert (r1.sum64, r1.xor32) == (r2.sum64, r2.xor32) == oc.vectoradd_digest(1 << 24, 8, 2)


def test_commutative_on_device(gp, torch_cuda):
    torch = torch_cuda
    n = 1 << 22
    da = torch.empty(n, device="cuda", dtype=torch.float32)
    db = torch.empty_like(da)
    c1 = torch.empty_like(da)
    c2 = torch.empty_like(da)
    s = torch.cuda.current_stream().cuda_stream
    gp.fill_inputs_device(da, db, n, seed=1, gen_kind=2, stream=s)
    gp.vectoradd_device(da, db, c1, n, stream=s)
    gp.vectoradd_device(db, da, c2, n, stream=s)
    torch.cuda.synchronize()
    assert torch.equal(c1.view(torch.int32), c2.view(torch.int32))


def test_sweep_2p30_against_oracle_digest(gp):
    """2^30 elements (12 GiB of traffic per launch): full digest against the oracle's streaming digest."""
    n = 1 << 30
    res = gp.vectoradd(n, iters=3, warmup=1, seed=0, gen_kind=0, verify=3)
    assert (res.sum64, res.xor32) == oc.vectoradd_digest(n, 0, 0)
    assert res.mismatches == 0 and res.l2_resident == 0


def test_full_size_2p32_properties(gp, torch_cuda):
    """BASELINE.json configs[1] upper end: 2^32 elements, 16 GiB per array, 64-bit indexing.
    Properties: (1) sampled slices equal the oracle bit for bit, (2) the digest of the whole
    equals the sum/xor of per-chunk digests (checksum of checksums), (3) the on-device
    regenerate-and-compare finds nothing."""
    torch = torch_cuda
    n = 1 << 32
    free, _ = torch.cuda.mem_get_info()
    if free < 3 * 4 * n + (2 << 30):
        pytest.skip("not enough free HBM for 3 x 16 GiB")
    s = torch.cuda.current_stream().cuda_stream
    da = torch.empty(n, device="cuda", dtype=torch.float32)
    db = torch.empty(n, device="cuda", dtype=torch.float32)
    dc = torch.empty(n, device="cuda", dtype=torch.float32)
    gp.fill_inputs_device(da, db, n, seed=0, gen_kind=0, stream=s)
    gp.vectoradd_device(da, db, dc, n, stream=s)
    torch.cuda.synchronize()
    m = 1 << 16
    for first in [0, (1 << 31) - m // 2, (1 << 32) - m, 3_000_000_019]:
        got = dc[first:first + m].cpu().numpy()
        a, b = oc.fill_inputs(m, 0, 0, first)
        assert oc.count_mismatch(got, oc.vectoradd(a, b)) == 0, first
    whole = gp.digest_device(dc, n, stream=s)
    parts_s, parts_x = 0, 0
    step = 1 << 29
    for lo in range(0, n, step):
        ps, px = gp.digest_device(dc.data_ptr() + 4 * lo, step, stream=s)
        parts_s = (parts_s + ps) & (2**64 - 1)
        parts_x ^= px
    assert whole == (parts_s, parts_x)
    # the first 2^28 elements of the big run also match the oracle's streaming digest
    assert gp.digest_device(dc, 1 << 28, stream=s) == oc.vectoradd_digest(1 << 28, 0, 0)
    assert gp.verify_device(dc, n, seed=0, gen_kind=0, stream=s) == 0
    del da, db, dc
    torch.cuda.empty_cache()


def test_error_paths_report_codes_not_crashes(gp, torch_cuda):
    """Out-of-memory, null pointers and a bad device are reported through the status code and message."""
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd(1 << 37, iters=1, warmup=0)  # 3 x 512 GiB
    assert e.value.code == gp.ENOMEM and "cudaMalloc" in e.value.message
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd_device(0, 0, 0, 16)
    assert e.value.code == gp.EINVAL
    with pytest.raises(gp.GpuProbeError) as e:
        gp.vectoradd(16, dev=63)
    assert e.value.code == gp.ENODEV
    with pytest.raises(gp.GpuProbeError) as e:
        gp.digest_device(0, 16)
    assert e.value.code == gp.EINVAL
    # the device is still usable afterwards
    assert gp.vectoradd(50_000, iters=1, warmup=0).mismatches == 0


def test_concurrent_callers_on_one_device_are_serialised(gp):
    """The header promises: safe from any OS thread, calls on one device serialised inside the library
    (a Go validator calls from whichever thread its goroutine happens to run on)."""
    import threading

    results, errors = {}, []

    def work(i):
        try:
            n = (1 << 22) + 1000 * i
            r = gp.vectoradd(n, iters=3, warmup=1, seed=i, gen_kind=i % 3)
            results[i] = (n, i, i % 3, r.sum64, r.xor32, r.mismatches)
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(i,)) for i in range(6)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for n, seed, kind, s, x, bad in results.values():
        assert bad == 0 and (s, x) == oc.vectoradd_digest(n, seed, kind)


def test_cached_staging_survives_growing_shrinking_and_shutdown(gp, torch_cuda, monkeypatch):
    """The staging slots and pinned bounce buffers of gpuprobe_vectoradd_buffers are cached per device and only grow;
    gpuprobe_shutdown frees them and the next call rebuilds them; GPUPROBE_BUF_CACHE=0 frees them after every call.
    Every result is checked against the host loop, and the cache must not leak: device memory in use after the calls
    is bounded by the largest staging set, and returns to the starting point after shutdown."""
    torch = torch_cuda
    torch.cuda.synchronize()
    gp.shutdown()
    free0, _ = torch.cuda.mem_get_info()
    sizes = [50_000, 7, (9 << 20) + 5, 50_000, 1 << 20, (9 << 20) + 5, 3]
    for rep, n in enumerate(sizes * 2):
        a, b = oc.fill_inputs(n, seed=20 + rep, kind=rep % 3)
        c = np.full(n, np.nan, dtype=np.float32)
        gp.vectoradd_buffers(a, b, c, n)
        assert oc.count_mismatch(c, oc.vectoradd(a, b)) == 0, (rep, n)
        if rep == len(sizes) - 1:
            gp.shutdown()  # mid-way: everything is rebuilt on demand
    free1, _ = torch.cuda.mem_get_info()
    assert free0 - free1 <= 3 * 3 * (8 << 20) * 4 + (64 << 20), "more device memory held than three slots of three 32 MiB chunks"
    gp.shutdown()
    free2, _ = torch.cuda.mem_get_info()
    assert free2 >= free0 - (8 << 20)
    monkeypatch.setenv("GPUPROBE_BUF_CACHE", "0")
    for n in (50_000, (8 << 20) + 1):
        a, b = oc.fill_inputs(n, seed=3, kind=0)
        c = np.empty(n, dtype=np.float32)
        gp.vectoradd_buffers(a, b, c, n)
        assert oc.count_mismatch(c, oc.vectoradd(a, b)) == 0
        free3, _ = torch.cuda.mem_get_info()
        assert free3 >= free0 - (8 << 20), "GPUPROBE_BUF_CACHE=0 must give the staging back after the call"


def test_digest_and_verify_share_cached_scratch_across_threads_and_streams(gp, torch_cuda):
    """digest/verify use one cached scratch buffer per device; callers on different host threads and streams are
    serialised on it inside the library and must all get the oracle's answer."""
    import threading

    torch = torch_cuda
    n = (1 << 22) + 77
    bufs, want = [], []
    for i in range(4):
        w = torch.empty(n, device="cuda", dtype=torch.float32)
        gp.fill_inputs_device(w, None, n, seed=40 + i, gen_kind=gp.GEN_BITS)
        bufs.append(w)
        a, _ = oc.fill_inputs(n, seed=40 + i, kind=oc.GEN_BITS)
        want.append(oc.digest(a))
    torch.cuda.synchronize()
    got, errors = {}, []

    def work(i):
        try:
            s = torch.cuda.Stream()
            for _ in range(20):
                got[i] = gp.digest_device(bufs[i], n, stream=s.cuda_stream)
                assert got[i] == want[i]
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


@pytest.mark.parametrize("kernel", [0, 1])
def test_p2p_kernels_loop_back_on_one_gpu_against_oracle(gp, kernel):
    """gpuprobe_p2p_loopback: the P2P copy+checksum kernels (0 = cp.async.bulk ring, 1 = register path) with both buffers
    on one device, so that their checksum reduction and the landed copy are checked against the oracle where only one GPU
    is visible. Sizes: below one 32 KiB chunk, ragged past several chunks, and large enough for every CTA to loop."""
    for nbytes, seed in [(16, 1), (32768 - 16, 2), ((3 << 20) + 16 * 5, 3), (256 << 20, 4)]:
        c = gp.p2p_loopback(nbytes, dev=0, iters=2, kernel=kernel, seed=seed)
        want = oc.p2p_checksum(nbytes // 4, seed, 0)
        assert c.verified == 1 and (c.src, c.dst) == (0, 0)
        assert (c.checksum_got, c.xor_got) == want == (c.checksum_expected, c.xor_expected)
    assert c.gbps > 500, f"{c.gbps:.0f} GB/s is not an HBM copy"
    with pytest.raises(gp.GpuProbeError) as e:
        gp.p2p_loopback(1000, kernel=kernel)  # not a multiple of 16
    assert e.value.code == gp.EINVAL
