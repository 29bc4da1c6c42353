"""Test helper for the raw-frame-stream entry point (zflacb200_open_frames): the oracle only speaks whole FLAC files, so a
raw frame stream is put behind a synthetic "fLaC" + STREAMINFO that states no total and no MD5 — what the reference
itself would be handed if somebody re-wrapped those frames (src/zflac.zig:224-240 for the field layout)."""
RATES = {1: 88200, 2: 176400, 3: 192000, 4: 8000, 5: 16000, 6: 22050, 7: 24000, 8: 32000, 9: 44100, 10: 48000, 11: 96000}
DEPTHS = {1: 8, 2: 12, 4: 16, 5: 20, 6: 24, 7: 32}
CHANNELS = [1, 2, 3, 4, 5, 6, 7, 8, 2, 2, 2]


def first_frame_params(frames: bytes, rate_hint: int = 0):
    """(channels, depth, sample_rate) from the 4 fixed header bytes of the frame `frames` starts with (zflac.zig:343-374)."""
    assert frames[0] == 0xFF and (frames[1] & 0xFE) == 0xF8
    sr_code, ch_code, bd_code = frames[2] & 15, frames[3] >> 4, (frames[3] >> 1) & 7
    return CHANNELS[ch_code], DEPTHS.get(bd_code, 0), RATES.get(sr_code, rate_hint if sr_code else 0)


def wrap_frames(frames: bytes, channels: int, depth: int, sample_rate: int) -> bytes:
    """fLaC + one last-block STREAMINFO (total = 0, MD5 = 0) + the frames."""
    si = bytearray(34)
    si[0:2] = (16).to_bytes(2, "big")
    si[2:4] = (65535).to_bytes(2, "big")
    packed = (sample_rate << 44) | ((channels - 1) << 41) | ((depth - 1) << 36)  # 20 + 3 + 5 + 36 bits
    si[10:18] = packed.to_bytes(8, "big")
    return b"fLaC" + bytes([0x80, 0, 0, 34]) + bytes(si) + frames
