"""gpuprobe_digest_device_noalloc -- the digest entry point for callers that cannot allocate.

Why it exists: a validator re-check may run while tenant jobs hold nearly all of the GPU's
memory, and cudaMalloc/cudaFree would also stall those jobs (device-wide locks, cudaFree waits
for the device to idle); callers inside a framework's memory pool may not use the driver
allocator at all.  This variant keeps its scratch in a fixed array inside the library, so it
has a size limit and all calls on a GPU share (and re-use) the same scratch.

The test calls it through the C ABI and requires the same sum64/xor32 as the allocating entry
point gpuprobe_digest_device, bit for bit (integer work: the tolerance is zero), on two calls in
a row -- the second with different data, so anything stale left in the shared scratch by the
first call would show.  It is collected last (file name) so that a -x run has already executed
every other test, and it must pass both in a fresh process and after the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available(), "this test needs a GPU"
    torch.cuda.set_device(0)
    return torch


def test_noalloc_digest_matches_allocating_digest_twice_in_a_row(gp, torch_cuda):
    torch = torch_cuda
    stream = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(2026)
    n1, n2 = (1 << 24) + 12_345, (1 << 22) + 3  # several segments each, ragged ends
    w1 = torch.from_numpy(rng.integers(0, 2**32, size=n1, dtype=np.uint32).view(np.int32)).cuda()
    w2 = torch.from_numpy(rng.integers(0, 2**32, size=n2, dtype=np.uint32).view(np.int32)).cuda()
    torch.cuda.synchronize()

    want1 = gp.digest_device(w1, n1, stream=stream)
    want2 = gp.digest_device(w2, n2, stream=stream)
    assert want1 != want2

    # the checker's own answer, computed on the CPU from the same words
    from oracle import oracle as oc

    host1 = w1.cpu().numpy().view(np.float32)
    assert want1 == oc.digest(host1)

    got1 = gp.digest_device_noalloc(w1, n1, stream=stream)  # first call on this device: does the set-up
    got2 = gp.digest_device_noalloc(w2, n2, stream=stream)  # second call re-uses the same scratch array
    assert got1 == want1
    assert got2 == want2
    # and again in the other order, smaller after larger and larger after smaller
    assert gp.digest_device_noalloc(w1, n1, stream=stream) == want1
    # unaligned start and tiny sizes go through the scalar path
    assert gp.digest_device_noalloc(w2.data_ptr() + 4, 1001, stream=stream) == gp.digest_device(w2.data_ptr() + 4, 1001, stream=stream)
    assert gp.digest_device_noalloc(w2, 1, stream=stream) == gp.digest_device(w2, 1, stream=stream)
    assert gp.digest_device_noalloc(w2, 0, stream=stream) == (0, 0)


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch(gp, torch_cuda):
    torch = torch_cuda
    w = torch.zeros(16, device="cuda", dtype=torch.int32)
    with pytest.raises(gp.GpuProbeError) as e:
        gp.digest_device_noalloc(w, gp.NOALLOC_MAX_WORDS + 1)  # rejected before anything is read
    assert e.value.code == gp.EINVAL
    assert "GPUPROBE_NOALLOC_MAX_WORDS" in e.value.message


def test_noalloc_at_its_maximum_size(gp, torch_cuda):
    torch = torch_cuda
    n = gp.NOALLOC_MAX_WORDS
    free, _ = torch.cuda.mem_get_info()
    if free < 4 * n + (1 << 30):
        pytest.skip("not enough free HBM for a 4 GiB buffer")
    w = torch.empty(n, device="cuda", dtype=torch.int32)
    stream = torch.cuda.current_stream().cuda_stream
    gp.fill_inputs_device(w, None, n, seed=5, gen_kind=gp.GEN_BITS, stream=stream)
    assert gp.digest_device_noalloc(w, n, stream=stream) == gp.digest_device(w, n, stream=stream)
