"""gp_pattern_verify_noalloc — the verify entry point for callers that cannot allocate.

Some users of this library hold device memory but may not create any: a probe that re-grades a
GPU while pods own nearly all of its memory, or code that must stay out of cudaMalloc/cudaFree
because they synchronise the device.  They hand in their own device buffer and a device slot
for the answer; the kernel's scratch lives in a fixed array inside the library.  This test
drives that entry point through the C ABI on memory torch allocated (standing in for the
caller's own), twice in a row on one stream, and requires the same bits as the ordinary
gp_pattern_host path on the same words.  The second call grades clean data right after a call
that saw faults, so anything the first call left in the shared scratch would show.

Collected last on purpose (see the file name): a run that stops at the first failure has
already run every other test.
"""
import numpy as np
import pytest

import gpuprobe

pytestmark = pytest.mark.gpu
M64 = (1 << 64) - 1


def _run(torch, host_words, kind, pas):
    d_words = torch.from_numpy(host_words.view(np.int64)).cuda()
    d_res = torch.full((6,), -1, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    gpuprobe.pattern_verify_noalloc(0, d_words.data_ptr(), host_words.size, kind, pas,
                                    d_res.data_ptr(), stream)
    torch.cuda.synchronize()
    return tuple(int(v) & M64 for v in d_res.cpu().numpy().view(np.uint64))


def test_noalloc_matches_existing_entry_point_twice_in_a_row(probe, oracle):
    import torch
    torch.cuda.set_device(0)
    n = 64 * 5000 + 64  # not a multiple of the 4096-word chunk: exercises the ragged last CTA
    i = np.arange(n, dtype=np.uint64)

    # call 1: walking-ones pass 9 with two faults planted
    words1 = np.left_shift(np.uint64(1), (i + np.uint64(9)) & np.uint64(63))
    words1[777] ^= np.uint64(1 << 40)
    words1[123456] ^= np.uint64(1)
    got1 = _run(torch, words1, gpuprobe.GP_PAT_WALKING_ONES, 9)
    # call 2, straight after: clean address-in-address pass 1
    words2 = ~(i * np.uint64(8))
    got2 = _run(torch, words2, gpuprobe.GP_PAT_ADDRESS, 1)

    ref1 = probe.pattern_host("0", words1, gpuprobe.GP_PAT_WALKING_ONES, 9, fill=False).astuple()
    ref2 = probe.pattern_host("0", words2, gpuprobe.GP_PAT_ADDRESS, 1, fill=False).astuple()
    assert got1 == ref1  # bit-exact, as every integer test in this suite
    assert got2 == ref2
    # and both against the CPU oracle on the same words, not only against the other CUDA entry point
    import oracle_lib
    assert got1 == oracle_lib.verify(oracle, words1, gpuprobe.GP_PAT_WALKING_ONES, 9).astuple()
    assert got2 == oracle_lib.verify(oracle, words2, gpuprobe.GP_PAT_ADDRESS, 1).astuple()
    assert got1[3:] == (2, 777, 1 << 40)
    assert got2[3:] == (0, M64, 0)


def test_noalloc_grades_a_sub_range_and_refuses_misaligned_pointers(oracle):
    """A caller grading part of its own buffer passes an interior pointer.  The kernels load 256 bits at
    a time, so the range has to start on a 32-byte boundary: an aligned interior pointer works, anything
    else is refused with GP_ERR_INVALID before a kernel could fault on it (a misaligned-address fault is
    sticky and would take the caller's whole context down)."""
    import torch
    import oracle_lib
    torch.cuda.set_device(0)
    n = 64 * 100
    i = np.arange(n, dtype=np.uint64)
    words = np.left_shift(np.uint64(1), (i + np.uint64(3)) & np.uint64(63))
    words[2000] ^= np.uint64(1 << 7)
    big = torch.zeros(n + 64, dtype=torch.int64, device="cuda")
    big[4:4 + n] = torch.from_numpy(words.view(np.int64)).cuda()       # +32 bytes: still aligned
    d_res = torch.full((8,), -1, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    gpuprobe.pattern_verify_noalloc(0, big.data_ptr() + 32, n, gpuprobe.GP_PAT_WALKING_ONES, 3,
                                    d_res.data_ptr() + 8, stream)      # result slot at +8: still aligned
    torch.cuda.synchronize()
    got = tuple(int(v) & M64 for v in d_res[1:7].cpu().numpy().view(np.uint64))
    assert got == oracle_lib.verify(oracle, words, gpuprobe.GP_PAT_WALKING_ONES, 3).astuple()
    assert got[3:] == (1, 2000, 1 << 7)
    for woff, roff in [(8, 0), (16, 0), (0, 4), (24, 4)]:
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            gpuprobe.pattern_verify_noalloc(0, big.data_ptr() + woff, 64, gpuprobe.GP_PAT_WALKING_ONES, 0,
                                            d_res.data_ptr() + roff, stream)
        assert e.value.code == gpuprobe.GP_ERR_INVALID and "aligned" in e.value.detail
    torch.cuda.synchronize()  # the refused calls launched nothing: the context is intact


def test_noalloc_rejects_oversize_and_bad_arguments():
    import torch
    d_res = torch.zeros(6, dtype=torch.int64, device="cuda")
    d_words = torch.zeros(64, dtype=torch.int64, device="cuda")
    with pytest.raises(gpuprobe.GpuProbeError) as e:
        gpuprobe.pattern_verify_noalloc(0, d_words.data_ptr(), gpuprobe.GP_NOALLOC_MAX_WORDS + 64,
                                        gpuprobe.GP_PAT_WALKING_ONES, 0, d_res.data_ptr(), 0)
    assert e.value.code == gpuprobe.GP_ERR_TOO_LARGE and "exceeds the limit" in e.value.detail
    for args in [(65, 0, 0), (64, 0, 64), (64, 7, 0)]:
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            gpuprobe.pattern_verify_noalloc(0, d_words.data_ptr(), args[0], args[1], args[2], d_res.data_ptr(), 0)
        assert e.value.code == gpuprobe.GP_ERR_INVALID
    with pytest.raises(gpuprobe.GpuProbeError) as e:
        gpuprobe.pattern_verify_noalloc(0, d_words.data_ptr(), 64, 0, 0, 0, 0)
    assert e.value.code == gpuprobe.GP_ERR_INVALID
