"""zflacb200_scan_no_alloc (include/zflacb200.h): the frame-header scan with its scratch in a static device array — no
counterpart in the reference, added for callers that may not call cudaMalloc. Through the C ABI, against the existing
entry points (zflacb200_open + zflacb200_frames) bit for bit: every frame of the decoded stream is a header the scan
reports, with the same offset / block size / header length / channel assignment, and what the scan reports beyond them
are chance matches that are no frame. Calls in a row share the array (the second call finds the first one's state in
it). This file sorts after every other test file on purpose."""
import numpy as np
import pytest

from tools import flacgen as fg

pytestmark = pytest.mark.gpu

HEADER = np.dtype([("offset", "<u8"), ("number", "<u8"), ("block_size", "<u4"), ("sample_rate", "<u4"), ("header_len", "u1"),
                   ("channel_code", "u1"), ("depth_code", "u1"), ("rate_code", "u1"), ("variable", "u1"), ("pad", "u1", (3,))])
CAP = 4096
MAX_BYTES = 1 << 30


def _lib():
    from zflac_b200 import _native
    return _native.load()


class Case:
    """A stream in device memory, its frames by the existing entry points, and output buffers — all allocated up front."""

    def __init__(self, torch, z, params):
        s = fg.generate(params)
        self.n = int(s.flac.size)
        self.d_stream = torch.zeros((self.n + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
        self.d_stream[: self.n] = torch.from_numpy(s.flac).cuda()
        h = z.open_device(0, 0, host_data=s.flac)
        try:
            self.frames = h.frames()
            self.audio_offset = int(h.info["audio_offset"])
            assert h.info["n_frames"] == len(self.frames) == s.n_frames
        finally:
            h.close()
        self.d_headers = torch.zeros(CAP * HEADER.itemsize, dtype=torch.uint8, device="cuda")
        self.d_count = torch.zeros(1, dtype=torch.int32, device="cuda")

    def scan(self, lib, stream=0, capacity=CAP):
        return lib.zflacb200_scan_no_alloc(self.d_stream.data_ptr(), self.n, self.audio_offset, self.d_headers.data_ptr(), capacity,
                                           self.d_count.data_ptr(), -1, stream)

    def result(self):
        count = int(self.d_count.cpu()[0])
        return count, self.d_headers.cpu().numpy().view(HEADER)[: min(count, CAP)].copy()

    def check(self, count, headers):
        frames = self.frames
        assert count <= CAP and len(headers) == count
        assert np.all(np.diff(headers["offset"].astype(np.int64)) > 0), "headers leave the scan in stream order"
        is_frame = np.isin(headers["offset"], frames["offset"])
        mine = headers[is_frame]
        assert len(mine) == len(frames), "every frame of the stream is a header the scan reports"
        for field_new, field_old in (("offset", "offset"), ("block_size", "block_size"), ("header_len", "header_len"),
                                     ("channel_code", "channel_code")):
            assert np.array_equal(mine[field_new], frames[field_old]), field_new
        extras = headers[~is_frame]  # chance matches inside frame data: valid grammar and CRC-8, but no frame starts there
        assert len(extras) <= 2 + len(frames) // 50, f"{len(extras)} headers that are no frames"
        return len(extras)


def test_two_calls_in_a_row_match_the_existing_entry_points():
    import torch
    import zflac_b200 as z
    lib = _lib()
    a = Case(torch, z, fg.c2(20.0))
    b = Case(torch, z, fg.c4_variable(10.0))  # another shape: variable block sizes, all channel assignments
    assert a.scan(lib) == 0, lib.zflacb200_last_error().decode()  # first call on this device: fetches the array's address
    torch.cuda.synchronize()
    first = a.result()
    a.check(*first)
    # the second call finds the first call's tile states and ticket in the array and must clear what it uses
    assert b.scan(lib) == 0, lib.zflacb200_last_error().decode()
    torch.cuda.synchronize()
    b.check(*b.result())
    # ... and back: the same input gives the same bytes
    a.d_headers.zero_()
    assert a.scan(lib) == 0
    torch.cuda.synchronize()
    again = a.result()
    assert again[0] == first[0] and again[1].tobytes() == first[1].tobytes()
    # two calls queued back to back on one stream, no synchronisation in between: stream order is enough. Device memory
    # is measured around these calls alone (the stream exists and has run work before, torch does nothing in between).
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        a.d_headers.zero_()
        b.d_headers.zero_()
    st.synchronize()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    assert a.scan(lib, st.cuda_stream) == 0 and b.scan(lib, st.cuda_stream) == 0
    st.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "the calls took device memory"
    a.check(*a.result())
    b.check(*b.result())


def test_limits_and_errors():
    import torch
    import zflac_b200 as z
    lib = _lib()
    a = Case(torch, z, fg.c2(5.0))
    # beyond the stated maximum: refused before anything is launched (the pointer is never touched)
    rc = lib.zflacb200_scan_no_alloc(a.d_stream.data_ptr(), MAX_BYTES + 17, 0, a.d_headers.data_ptr(), CAP, a.d_count.data_ptr(), -1, 0)
    assert rc == 101 and b"ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES" in lib.zflacb200_last_error()
    # a pointer that is not 16-byte aligned, a null pointer
    assert lib.zflacb200_scan_no_alloc(a.d_stream.data_ptr() + 1, a.n - 1, 0, a.d_headers.data_ptr(), CAP, a.d_count.data_ptr(), -1, 0) == 101
    assert b"aligned" in lib.zflacb200_last_error()
    assert lib.zflacb200_scan_no_alloc(0, a.n, 0, a.d_headers.data_ptr(), CAP, a.d_count.data_ptr(), -1, 0) == 101
    # fewer records than headers: the count says how many there were, the first `capacity` are stored
    assert a.scan(lib) == 0
    torch.cuda.synchronize()
    count, headers = a.result()
    a.check(count, headers)
    a.d_headers.zero_()
    assert a.scan(lib, capacity=5) == 0
    torch.cuda.synchronize()
    count5, _ = a.result()
    got = a.d_headers.cpu().numpy().view(HEADER)
    assert count5 == count and got[:5].tobytes() == headers[:5].tobytes() and not got[5:].tobytes().strip(b"\0")
    # an empty range: zero headers, no launch
    assert lib.zflacb200_scan_no_alloc(a.d_stream.data_ptr(), a.n, a.n, a.d_headers.data_ptr(), CAP, a.d_count.data_ptr(), -1, 0) == 0
    torch.cuda.synchronize()
    assert int(a.d_count.cpu()[0]) == 0
