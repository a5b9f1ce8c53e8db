"""Fabric probes on real hardware: all-pairs NVLink P2P copy+checksum and the NCCL int32
allreduce.  Need >= 2 GPUs (gpurun --gpus 2); skipped on a single-GPU box."""
import pytest

from oracle import oracle as oc

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


@pytest.fixture(scope="module")
def ndev(gp):
    n = gp.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    return n


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_p2p_matrix_checksums_match_oracle(gp, ndev, mode):
    nd = min(ndev, 8)
    nbytes = (64 << 20) + 16 * 7  # ragged: not a multiple of the kernel's tile
    rc, cells = gp.p2p_matrix(nd, nbytes, iters=3, warmup=1, mode=mode, seed=3)
    assert rc == gp.OK
    for s in range(nd):
        want = oc.p2p_checksum(nbytes // 4, 3, s)
        for t in range(nd):
            c = cells[s][t]
            assert (c.src, c.dst) == (s, t)
            if s == t:
                assert c.gbps == 0
                continue
            assert c.peer_ok == 1 and c.verified == 1
            assert (c.checksum_got, c.xor_got) == want == (c.checksum_expected, c.xor_expected)
            assert c.gbps > 50, f"{s}->{t}: {c.gbps:.1f} GB/s is not NVLink"


def test_p2p_small_and_explicit_device_list(gp, ndev):
    rc, cells = gp.p2p_matrix(2, 16, iters=1, warmup=0, seed=1, devs=[1, 0])
    assert rc == gp.OK
    assert (cells[0][1].src, cells[0][1].dst) == (1, 0)
    assert (cells[0][1].checksum_got, cells[0][1].xor_got) == oc.p2p_checksum(4, 1, 1)
    assert (cells[1][0].checksum_got, cells[1][0].xor_got) == oc.p2p_checksum(4, 1, 0)
    rc, _ = gp.p2p_matrix(2, 1 << 20, devs=[0, 0], raise_on_fail=False)
    assert rc == gp.EINVAL


def assert_reduced_buffer_matches_oracle(res, nd, count, seed, warmup, iters):
    """The collective's OUTPUT against the CPU oracle: the library returns sum64/xor32 of the buffer rank 0 holds
    after the last in-place round (and whether every other rank holds the same); the oracle computes the digest of
    nd^(rounds-1) * sum_r x_r[i] from its own generator.  `mismatches` is the library's self-check and is only
    looked at second."""
    assert res.ndev == nd and res.bytes == 4 * count
    assert res.rounds == max(1, warmup) + iters
    assert res.ranks_equal == 1
    assert (res.sum64, res.xor32) == oc.allreduce_digest(count, seed, nd, rounds=res.rounds)
    if count <= 1 << 16:  # and once more through the array form of the oracle, multiplied out in numpy
        import numpy as np

        e = oc.allreduce_expected(count, seed, nd).view(np.uint32) * np.uint32(pow(nd, res.rounds - 1, 1 << 32))
        assert res.sum64 == int(e.sum(dtype=np.uint64)) and res.xor32 == int(np.bitwise_xor.reduce(e))
    assert res.mismatches == 0


AR_COUNTS = [1, 5, 1000, 4096 * 3 + 5, (16 << 20) + 7]


@pytest.mark.parametrize("count", AR_COUNTS)
def test_allreduce_int32_bit_exact(gp, ndev, count):
    """NCCL int32 sum allreduce (north_star item c): reduced buffers against the oracle, bit for bit."""
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        res = gp.allreduce_i32(nd, count, iters=3, warmup=2, seed=2)
        assert_reduced_buffer_matches_oracle(res, nd, count, 2, 2, 3)
        assert res.nccl_version >= 20000 and res.impl == gp.AR_IMPL_NCCL


@pytest.mark.parametrize("count", AR_COUNTS)
def test_nvls_allreduce_kernel_bit_exact(gp, ndev, count):
    """The hand-written multimem kernel (fill -> ld_reduce/st): reduced buffers against the oracle."""
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        res = gp.allreduce_i32_nvls(nd, count, iters=3, warmup=2, seed=4)
        assert_reduced_buffer_matches_oracle(res, nd, count, 4, 2, 3)
        assert res.nccl_version == 0 and res.impl == gp.AR_IMPL_NVLS


def test_nvls_allreduce_with_16_byte_multicast_stores(gp, ndev, monkeypatch):
    """GPUPROBE_NVLS_ST_V4=1: the shuffle-transposed 16-byte multimem.st path, every unroll the library builds."""
    monkeypatch.setenv("GPUPROBE_NVLS_ST_V4", "1")
    for unroll in ("4", "8", "16"):
        monkeypatch.setenv("GPUPROBE_NVLS_UNROLL", unroll)
        for nd in [n for n in (2, 4, 8) if n <= ndev]:
            for count in (1000, (4 << 20) + 3):
                res = gp.allreduce_i32_nvls(nd, count, iters=2, warmup=1, seed=5)
                assert_reduced_buffer_matches_oracle(res, nd, count, 5, 1, 2)
                res = gp.allreduce_i32_hybrid(nd, count, iters=2, warmup=1, seed=5)
                assert_reduced_buffer_matches_oracle(res, nd, count, 5, 1, 2)


def test_nvls_and_nccl_agree_on_bandwidth_order_of_magnitude(gp, ndev):
    nd = 8 if ndev >= 8 else (4 if ndev >= 4 else 2)
    a = gp.allreduce_i32_nvls(nd, 1 << 26, iters=5, warmup=2)
    b = gp.allreduce_i32(nd, 1 << 26, iters=5, warmup=2)
    assert a.mismatches == 0 and b.mismatches == 0
    assert (a.sum64, a.xor32) == (b.sum64, b.xor32) == oc.allreduce_digest(1 << 26, 0, nd, rounds=7)
    assert a.busbw_gbps > 0.3 * b.busbw_gbps


@pytest.mark.parametrize("count", AR_COUNTS)
def test_p2p_two_shot_allreduce_kernel_bit_exact(gp, ndev, count):
    """Same buffers and barriers as the NVLS kernel, data over plain peer mappings."""
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        res = gp.allreduce_i32_p2p(nd, count, iters=3, warmup=2, seed=6)
        assert_reduced_buffer_matches_oracle(res, nd, count, 6, 2, 3)
        assert res.impl == gp.AR_IMPL_P2P
    with pytest.raises(gp.GpuProbeError) as e:
        gp.allreduce_i32_p2p(3, 1024) if ndev >= 3 else gp.allreduce_i32_p2p(1, 1024)
    assert e.value.code == gp.EINVAL


@pytest.mark.parametrize("count", AR_COUNTS)
def test_hybrid_allreduce_kernel_bit_exact(gp, ndev, count):
    """NVLS head + peer-mapping tail on disjoint CTAs of one kernel."""
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        res = gp.allreduce_i32_hybrid(nd, count, iters=3, warmup=2, seed=8)
        assert_reduced_buffer_matches_oracle(res, nd, count, 8, 2, 3)
        assert res.impl == gp.AR_IMPL_HYBRID


@pytest.mark.parametrize("count", AR_COUNTS + [31, 32, 33, 64 * 8 + 1])
def test_push_two_shot_allreduce_kernel_bit_exact(gp, ndev, count):
    """Posted writes only: scatter into the owner's staging area, local add, broadcast.  Counts around the 32-word
    slice rounding leave some ranks with an empty slice."""
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        res = gp.allreduce_i32_push(nd, count, iters=3, warmup=2, seed=10)
        assert_reduced_buffer_matches_oracle(res, nd, count, 10, 2, 3)
        assert res.impl == gp.AR_IMPL_PUSH


@pytest.mark.parametrize("count", AR_COUNTS + [31, 33, (1 << 21) + 1, (3 << 20) + 64])
def test_copy_engine_two_shot_allreduce_bit_exact(gp, ndev, count, monkeypatch):
    """Copy engines move every byte (cudaMemcpyPeerAsync ordered by events across the GPUs' streams), a kernel adds;
    chunk counts from one to more than the default, sizes that leave empty slices and ragged last chunks."""
    for chunks in ("1", "4", "16"):
        monkeypatch.setenv("GPUPROBE_CE_CHUNKS", chunks)
        for nd in [n for n in (2, 3, 4, 8) if n <= ndev]:
            res = gp.allreduce_i32_ce(nd, count, iters=3, warmup=2, seed=12)
            assert_reduced_buffer_matches_oracle(res, nd, count, 12, 2, 3)
            assert res.impl == gp.AR_IMPL_CE


@pytest.mark.parametrize("count", [1, 1000, 4096 * 3 + 5, (4 << 20) + 7, (16 << 20) + 7])
def test_nvls_plus_copy_engine_allreduce_bit_exact(gp, ndev, count, monkeypatch):
    """The multimem kernel reduces the head of the buffer while the copy-engine two-shot reduces the tail; splits from
    almost-all-switch to almost-all-engines, counts so small that the tail (or a slice of it) is empty."""
    for permille in ("800", "100", "999"):
        monkeypatch.setenv("GPUPROBE_NVLSCE_PERMILLE", permille)
        for nd in [n for n in (2, 3, 4, 8) if n <= ndev]:
            res = gp.allreduce_i32_nvls_ce(nd, count, iters=3, warmup=2, seed=13)
            assert_reduced_buffer_matches_oracle(res, nd, count, 13, 2, 3)
            assert res.impl == gp.AR_IMPL_NVLS_CE


@pytest.mark.parametrize("mode", [4, 5])
def test_p2p_copy_engine_modes_land_the_oracles_words(gp, ndev, mode):
    """SERIAL_CE / ROUNDS_CE: the destination's copy engine moves the bytes; the checksum is the digest of the landed copy."""
    nd = min(ndev, 8)
    nbytes = (32 << 20) + 16 * 3
    rc, cells = gp.p2p_matrix(nd, nbytes, iters=2, warmup=1, mode=mode, seed=5)
    assert rc == gp.OK
    for s in range(nd):
        want = oc.p2p_checksum(nbytes // 4, 5, s)
        for t in range(nd):
            if s != t:
                c = cells[s][t]
                assert c.verified == 1 and (c.checksum_got, c.xor_got) == want and c.gbps > 50


def test_auto_allreduce_picks_a_kernel_and_matches_oracle(gp, ndev):
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        res = gp.allreduce_i32_auto(nd, (8 << 20) + 9, iters=2, warmup=1, seed=11)  # 32 MiB: engines at 2 ranks only
        assert_reduced_buffer_matches_oracle(res, nd, (8 << 20) + 9, 11, 1, 2)
        assert res.impl == (gp.AR_IMPL_CE if nd == 2 else gp.AR_IMPL_HYBRID)
        res = gp.allreduce_i32_auto(nd, (1 << 20) + 9, iters=2, warmup=1, seed=11)  # 4 MiB: one cooperative kernel
        assert_reduced_buffer_matches_oracle(res, nd, (1 << 20) + 9, 11, 1, 2)
        assert res.impl == (gp.AR_IMPL_PUSH if nd == 2 else gp.AR_IMPL_HYBRID)
    if ndev >= 4:
        res = gp.allreduce_i32_auto(4, 1 << 27, iters=2, warmup=1, seed=11)  # 512 MiB at 4 ranks: engines
        assert_reduced_buffer_matches_oracle(res, 4, 1 << 27, 11, 1, 2)
        assert res.impl == gp.AR_IMPL_CE


def test_two_shot_allreduce_without_multicast(gp, ndev, monkeypatch):
    """GPUPROBE_NO_MULTICAST=1: plain peer mappings and a peer-store barrier (boxes without NVLS)."""
    monkeypatch.setenv("GPUPROBE_NO_MULTICAST", "1")
    for nd in [n for n in (2, 4, 8) if n <= ndev]:
        for count in (5, 4096 * 3 + 5, (4 << 20) + 1):
            res = gp.allreduce_i32_p2p(nd, count, iters=3, warmup=2, seed=9)
            assert_reduced_buffer_matches_oracle(res, nd, count, 9, 2, 3)
            res = gp.allreduce_i32_push(nd, count, iters=3, warmup=2, seed=9)
            assert_reduced_buffer_matches_oracle(res, nd, count, 9, 2, 3)
            res = gp.allreduce_i32_auto(nd, count, iters=2, warmup=1, seed=9)  # 2 ranks: push; else falls back to two-shot
            assert_reduced_buffer_matches_oracle(res, nd, count, 9, 1, 2)
            assert res.impl == (gp.AR_IMPL_PUSH if nd == 2 else gp.AR_IMPL_P2P)
    with pytest.raises(gp.GpuProbeError) as e:
        gp.allreduce_i32_nvls(2, 1024, iters=1, warmup=1)
    assert e.value.code == gp.ENOPEER
