"""GPU: vq_nms1d_segmented_static, the entry point that allocates nothing (include/vqwild.h; csrc/vq_noalloc.cu: the
kernel's scratch is a file-scope device array of VQ_NMS1D_STATIC_MAX bytes), called through the C-ABI and checked
against vq_nms1d_segmented on the same detections -- keep masks are bytes, so the bar is equality.  Collected last on
purpose; runs alone in a fresh process as well as after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX = 1048576  # VQ_NMS1D_STATIC_MAX


def _detections(seed, seg_len):
    """Moment-like (start, end, score) rows per segment; scores quantised so that ties inside a segment are common."""
    rng = np.random.RandomState(seed)
    seg_len = np.asarray(seg_len, np.int64)
    off = np.concatenate([[0], np.cumsum(seg_len)]).astype(np.int64)
    total = int(off[-1])
    start = np.floor(rng.uniform(0, 200, total))
    length = rng.choice([4.0, 8.0, 16.0, 32.0, 64.0], total)
    score = np.round(rng.uniform(-4, 0, total), 2)
    dets = np.stack([start, start + length, score], 1).astype(np.float64)
    return np.ascontiguousarray(dets), off


def _existing(L, dets, off, thresh):
    keep = np.zeros(dets.shape[0], np.uint8)
    L.check(L.lib().vq_nms1d_segmented(L.ptr(dets), L.ptr(off), C.c_int64(len(off) - 1), C.c_double(thresh), L.ptr(keep),
                                       C.c_void_p(0)))
    return keep


def _static(L, torch, dets, off, thresh, total=None):
    d = torch.from_numpy(dets).cuda()
    o = torch.from_numpy(off).cuda()
    keep = torch.full((max(dets.shape[0], 1),), 7, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = L.lib().vq_nms1d_segmented_static(L.ptr(d), L.ptr(o), C.c_int64(len(off) - 1),
                                           C.c_int64(dets.shape[0] if total is None else total), C.c_double(thresh),
                                           L.ptr(keep), L.stream_of(d))
    L.check(rc)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0, "the no-alloc entry point took device memory"
    return keep[:dets.shape[0]].cpu().numpy()


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from vqwild_b200 import _lib as L
    torch.cuda.set_device(0)
    warm, woff = _detections(9, [3])
    _existing(L, warm, woff, 0.5)  # loads nms1d_kernel's module, so that the memory check below sees only the call
    # set-up call: resolves the scratch array's device address on this device, launches nothing
    L.check(L.lib().vq_nms1d_segmented_static(None, None, C.c_int64(0), C.c_int64(0), C.c_double(0.5), None, C.c_void_p(0)))
    return L, torch


def test_static_nms_equals_existing_entry_point_two_calls_in_a_row(env):
    L, torch = env
    rng = np.random.RandomState(0)
    # call 1: ragged segments incl. empty ones, singletons and one as long as the longest real video (3 523 moments)
    len_a = np.concatenate([[0, 1, 3523, 0, 2, 33], rng.randint(0, 400, 300)])
    # call 2: fewer, differently cut rows -- what call 1 left in the shared array lies inside and beyond its range
    len_b = np.concatenate([[64, 0, 1, 1000], rng.randint(1, 90, 150)])
    a, off_a = _detections(1, len_a)
    b, off_b = _detections(2, len_b)
    want_a, want_b = _existing(L, a, off_a, 0.5), _existing(L, b, off_b, 0.3)
    got_a = _static(L, torch, a, off_a, 0.5)
    got_b = _static(L, torch, b, off_b, 0.3)
    assert np.array_equal(got_a, want_a)
    assert np.array_equal(got_b, want_b)
    assert want_a.sum() > len(len_a) // 2 and want_a.sum() < len(want_a)  # the case suppresses and keeps
    # and again, in the other order, on the contents the previous calls left behind
    assert np.array_equal(_static(L, torch, b, off_b, 0.3), want_b)
    assert np.array_equal(_static(L, torch, a, off_a, 0.5), want_a)


def test_static_nms_matches_oracle_on_a_few_segments(env):
    from oracle import oracle as orc
    L, torch = env
    dets, off = _detections(3, [40, 0, 7, 120])
    got = _static(L, torch, dets, off, 0.5)
    import vqwild_b200 as vq
    mirror = vq.nms_segmented_static(torch.from_numpy(dets).cuda(), torch.from_numpy(off).cuda(), 0.5)
    assert np.array_equal(mirror.cpu().numpy(), got)
    for s in range(len(off) - 1):
        want = np.zeros(off[s + 1] - off[s], np.uint8)
        want[orc.nms1d(dets[off[s]:off[s + 1]], 0.5)] = 1
        assert np.array_equal(got[off[s]:off[s + 1]], want)


def test_static_nms_at_the_maximum_and_beyond(env):
    L, torch = env
    rng = np.random.RandomState(4)
    lens = []
    left = STATIC_MAX
    while left:
        n = min(left, int(rng.randint(1, 200)))
        lens.append(n)
        left -= n
    dets, off = _detections(5, lens)
    assert dets.shape[0] == STATIC_MAX
    assert np.array_equal(_static(L, torch, dets, off, 0.5), _existing(L, dets, off, 0.5))
    # one detection more than the array holds: refused before anything is launched, with the limit in the message
    d = torch.zeros(8, 3, dtype=torch.float64, device="cuda")
    o = torch.zeros(2, dtype=torch.int64, device="cuda")
    k = torch.zeros(8, dtype=torch.uint8, device="cuda")
    rc = L.lib().vq_nms1d_segmented_static(L.ptr(d), L.ptr(o), C.c_int64(1), C.c_int64(STATIC_MAX + 1), C.c_double(0.5),
                                           L.ptr(k), C.c_void_p(0))
    assert rc == -6 and b"VQ_NMS1D_STATIC_MAX" in L.lib().vq_last_error()
    # host memory would have to be staged through an allocation: refused
    hd, ho = _detections(6, [5])
    hk = np.zeros(5, np.uint8)
    rc = L.lib().vq_nms1d_segmented_static(L.ptr(hd), L.ptr(ho), C.c_int64(1), C.c_int64(5), C.c_double(0.5), L.ptr(hk),
                                           C.c_void_p(0))
    assert rc == -1 and b"device memory only" in L.lib().vq_last_error()
