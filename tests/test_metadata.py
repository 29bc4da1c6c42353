"""CPU (host-only entry points, no device): zflacb200_metadata / _seektable / _vorbis_comments against
an independent walk in Python and against the oracle — same errors as the reference's metadata walk
(src/zflac.zig:217-253, :309), same audio offset, and every seek point lands on a frame of the oracle's chain."""
import struct

import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries, corpus_ids

HEAD_ERRORS = ("InvalidSignature", "InvalidMetadataHeader", "MissingStreaminfo")


def py_walk(data: bytes):
    assert data[:4] == b"fLaC"
    pos, blocks = 4, []
    while True:
        t, last = data[pos] & 0x7F, data[pos] >> 7
        length = int.from_bytes(data[pos + 1: pos + 4], "big")
        stored = 34 if t == 0 else length
        blocks.append((t, pos + 4, length, stored, last))
        pos += 4 + stored
        if last:
            return blocks, pos


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


@pytest.mark.parametrize("entry", corpus_entries(), ids=corpus_ids())
def test_corpus_metadata(z, entry):
    from oracle import oracle
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True, keep_on_md5_mismatch=True, copy_samples=False)
    if ref.error in HEAD_ERRORS:
        with pytest.raises(z.ZflacError) as ei:
            z.metadata(data)
        assert type(ei.value).__name__ == ref.error
        return
    m = z.metadata(data)
    blocks, audio = py_walk(data)
    assert m["audio_offset"] == audio == ref.audio_offset
    assert [(z.BLOCK_TYPES.index(b["type"]), b["offset"], b["length"], b["stored_bytes"], int(b["is_last"])) for b in m["blocks"]] == blocks
    # SEEKTABLE: every real point is a frame of the oracle's chain, at that sample number
    st = [b for b in blocks if b[0] == 3]
    assert len(m["seektable"]) == (st[0][3] // 18 if st else 0)
    if st and ref.frames is not None and len(ref.frames):
        ch = max(ref.channels, 1)
        starts = {int(o): int(s) // ch for o, s in zip(ref.frames["offset"], ref.frames["sample_offset"])}
        for k, p in enumerate(m["seektable"]):
            sn, so, fs = struct.unpack(">QQH", data[st[0][1] + 18 * k: st[0][1] + 18 * k + 18])
            assert (p["sample_number"], p["stream_offset"], p["frame_samples"]) == (sn, so, fs)
            if p["is_placeholder"]:
                continue
            at = audio + p["stream_offset"]
            if at in starts:  # (a truncated or faulty file may list points past what decodes)
                assert starts[at] == p["sample_number"]
            else:
                assert entry["expect_error"] or at >= int(ref.frames["end_offset"][-1])
    # VORBIS_COMMENT
    vc = [b for b in blocks if b[0] == 4]
    if (entry["group"], entry["id"]) == ("faulty", "10"):  # "invalid vorbis comment metadata block": decodes, comments do not parse
        assert m["comments_error"] == "InvalidMetadataHeader" and m["comments"] == [] and not entry["expect_error"]
    elif vc:
        o = vc[0][1]
        vl = int.from_bytes(data[o: o + 4], "little")
        assert m["vendor"] == data[o + 4: o + 4 + vl].decode("utf-8", "replace")
        n = int.from_bytes(data[o + 4 + vl: o + 8 + vl], "little")
        assert len(m["comments"]) == n
        q = o + 8 + vl
        for c in m["comments"][:50]:
            cl = int.from_bytes(data[q: q + 4], "little")
            assert c == data[q + 4: q + 4 + cl].decode("utf-8", "replace")
            q += 4 + cl
    else:
        assert m["vendor"] == "" and m["comments"] == []


def test_head_errors_in_the_reference_order(z):
    si = bytes([0x80, 0, 0, 34]) + bytes(34)
    with pytest.raises(z.InvalidSignature):
        z.metadata(b"fLaX" + si)
    with pytest.raises(z.EndOfStream):
        z.metadata(b"fL")
    with pytest.raises(z.EndOfStream):
        z.metadata(b"fLaC" + si[:20])
    with pytest.raises(z.InvalidMetadataHeader):
        z.metadata(b"fLaC" + bytes([0x07, 0, 0, 0]) + si)
    with pytest.raises(z.MissingStreaminfo):
        z.metadata(b"fLaC" + bytes([0x81, 0, 0, 4]) + bytes(4))
    with pytest.raises(z.EndOfStream):  # a block longer than the data
        z.metadata(b"fLaC" + bytes([0x00, 0, 0, 34]) + bytes(34) + bytes([0x81, 0, 1, 0]) + bytes(10))
    # comment lengths overrun the block: the walk is fine (the reference skips the block), the comments are not
    m = z.metadata(b"fLaC" + bytes([0x00, 0, 0, 34]) + bytes(34) + bytes([0x84, 0, 0, 8]) + (100).to_bytes(4, "little") + bytes(4))
    assert m["comments_error"] == "InvalidMetadataHeader" and m["audio_offset"] == 54
    m = z.metadata(b"fLaC" + bytes([0x80, 0, 0, 99]) + bytes(34))  # Q13: STREAMINFO is 34 bytes whatever its length field says
    assert m["audio_offset"] == 42 and m["blocks"][0]["length"] == 99 and m["blocks"][0]["stored_bytes"] == 34


def _py_pictures(data: bytes):
    blocks, _ = py_walk(data)
    out = []
    for t, off, _length, stored, _last in blocks:
        if t != 6:
            continue
        p = off
        ptype, mlen = struct.unpack(">II", data[p: p + 8]); p += 8
        mime = data[p: p + mlen]; p += mlen
        (dlen,) = struct.unpack(">I", data[p: p + 4]); p += 4
        desc = data[p: p + dlen]; p += dlen
        w, h, d, c, ilen = struct.unpack(">IIIII", data[p: p + 20]); p += 20
        assert p + ilen <= off + stored
        out.append((ptype, mime, desc, w, h, d, c, p, ilen))
    return out


@pytest.mark.parametrize("entry", [e for e in corpus_entries() if not e["expect_error"]],
                         ids=[i for i, e in zip(corpus_ids(), corpus_entries()) if not e["expect_error"]])
def test_pictures_and_cuesheet(z, entry):
    """PICTURE (subset 56-59: JPG / PNG / GIF / AVIF) and CUESHEET against an independent walk; the image bytes carry
    their format's magic number, so the spans point at real pictures."""
    data = corpus_bytes(entry)
    want = _py_pictures(data)
    got = z.pictures(data)
    assert len(got) == len(want)
    for g, (ptype, mime, desc, w, h, d, c, ioff, ilen) in zip(got, want):
        assert (g["picture_type"], g["width"], g["height"], g["depth"], g["colors"]) == (ptype, w, h, d, c)
        assert g["mime"].encode() == mime and g["description"].encode("utf-8") == desc and g["image"] == (ioff, ilen)
        magic = {"image/jpeg": b"\xff\xd8\xff", "image/png": b"\x89PNG", "image/gif": b"GIF8"}.get(g["mime"])
        if magic:
            assert data[ioff: ioff + len(magic)] == magic
    if entry["group"] == "subset" and entry["id"] in ("56", "57", "58", "59"):
        assert len(got) >= 1
    blocks, _ = py_walk(data)
    sheet = z.cuesheet(data)
    assert (sheet is not None) == any(b[0] == 5 for b in blocks)


def test_cuesheet_synthetic(z):
    """No corpus file available here carries a CUESHEET: build one (RFC 9639 8.7) in front of a real stream."""
    e = next(x for x in corpus_entries() if (x["group"], x["id"]) == ("subset", "01"))
    data = corpus_bytes(e)
    tracks = [(0, 1, b"USRC17607839", 0x00, [(0, 1)]), (441000, 2, b"\0" * 12, 0x40, [(0, 0), (588, 1)]), (882000, 170, b"\0" * 12, 0x00, [])]
    body = b"1234567890123".ljust(128, b"\0") + struct.pack(">Q", 88200) + bytes([0x80]) + bytes(258) + bytes([len(tracks)])
    for off, num, isrc, flags, idx in tracks:
        body += struct.pack(">QB", off, num) + isrc + bytes([flags]) + bytes(13) + bytes([len(idx)])
        for ioff, inum in idx:
            body += struct.pack(">QB", ioff, inum) + bytes(3)
    # insert after STREAMINFO (clear nothing: STREAMINFO is not the last block in subset/01)
    cue = bytes([5]) + len(body).to_bytes(3, "big") + body
    assert not data[4] >> 7
    new = data[:42] + cue + data[42:]
    sheet = z.cuesheet(new)
    assert sheet["media_catalog_number"] == "1234567890123" and sheet["is_cd"] and sheet["lead_in_samples"] == 88200
    assert [(t["offset_samples"], t["number"], t["is_audio"], t["pre_emphasis"]) for t in sheet["tracks"]] == [
        (0, 1, True, False), (441000, 2, True, True), (882000, 170, True, False)]
    assert sheet["tracks"][0]["isrc"] == "USRC17607839"
    assert sheet["tracks"][1]["indices"] == [dict(offset_samples=0, number=0), dict(offset_samples=588, number=1)]
    assert z.metadata(new)["audio_offset"] == z.metadata(data)["audio_offset"] + len(cue)
    with pytest.raises(z.InvalidMetadataHeader):       # track table overruns the block
        z.cuesheet(data[:42] + bytes([5]) + (400).to_bytes(3, "big") + body[:400] + data[42:])


def test_pcm_bound(z):
    from zflac_b200 import _native
    import ctypes as C
    lib = _native.load()
    for e in corpus_entries():
        data = corpus_bytes(e)
        arr = np.frombuffer(data, dtype=np.uint8)
        b = C.c_uint64(123)
        assert lib.zflacb200_pcm_bound(arr.ctypes.data, arr.size, C.byref(b)) == 0
        if e["expect_error"] in ("InvalidSignature", "InvalidMetadataHeader", "MissingStreaminfo"):
            assert b.value == 0
        elif not e["expect_error"] and b.value:
            assert b.value <= e["n_samples"] * e["container_bytes"]  # never more than what decodes (faulty/05 decodes more)


def _random_head(rng, stream: bytes) -> bytes:
    """`stream` = fLaC + one STREAMINFO + frames (tests/bitcraft.py). The head is rebuilt at random: block order, types
    (reserved and invalid ones too), lengths, last-block flags, a damaged signature, a cut anywhere in the head."""
    sig, si, frames = stream[:4], stream[8:42], stream[42:]
    blocks = []
    for _ in range(int(rng.integers(0, 5))):
        typ = int(rng.choice([1, 1, 2, 3, 4, 5, 6, 7, 50, 126, 127, 0]))
        n = int(rng.choice([0, 1, 4, 17, 34, 300]))
        blocks.append((typ, bytes(rng.integers(0, 256, n, dtype=np.uint8)) if typ else si))
    at = int(rng.integers(0, len(blocks) + 1)) if rng.random() < 0.2 else 0  # STREAMINFO first, mostly
    if rng.random() < 0.9:
        blocks.insert(at, (0, si))
    last = len(blocks) - 1 if rng.random() < 0.85 else int(rng.integers(-1, len(blocks)))  # -1: no block says it is the last
    head = bytearray(sig)
    for k, (typ, body) in enumerate(blocks):
        length = len(body) if rng.random() < 0.9 else int(rng.integers(0, 400))  # the length field may lie
        head += bytes([(0x80 if k == last else 0) | typ]) + length.to_bytes(3, "big") + body
    if rng.random() < 0.05:
        head[int(rng.integers(0, 4))] ^= 0x20
    data = bytes(head) + frames
    if rng.random() < 0.15:
        data = data[:int(rng.integers(0, len(head) + 8))]
    return data


def test_random_heads_against_the_oracle(z):
    """The host-only metadata walk (the one `decode()` starts with) on 600 random heads: where the oracle's decode fails in
    the head, the same error; where it gets past the head, the same audio offset."""
    import bitcraft
    from oracle import oracle
    rng = np.random.default_rng(0xFEED)
    base = [bitcraft.random_stream(seed)[0].tobytes() for seed in (3, 4, 5, 6)]
    seen = {}
    for k in range(600):
        data = _random_head(rng, base[k % len(base)])
        ref = oracle.decode(np.frombuffer(data, dtype=np.uint8), keep_on_md5_mismatch=True, copy_samples=False)
        try:
            m = z.metadata(data)
        except z.ZflacError as e:
            name = type(e).__name__
            assert ref.error == name, (k, name, ref.error, data[:64].hex())
            seen[name] = seen.get(name, 0) + 1
            continue
        seen["walked"] = seen.get("walked", 0) + 1
        # the oracle got past the head as well: whatever it says afterwards is about the frames
        assert ref.error not in ("InvalidSignature", "InvalidMetadataHeader", "MissingStreaminfo"), (k, ref.error, data[:64].hex())
        if ref.audio_offset:
            assert m["audio_offset"] == ref.audio_offset, (k, m["audio_offset"], ref.audio_offset)
    assert seen.get("walked", 0) > 100 and len(seen) >= 4, seen
