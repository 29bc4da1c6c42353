"""Hand-made FLAC frames for cases no encoder produces (tests only): an MSB-first bit writer, the two CRCs and a frame of
one-channel LPC subframes whose fields are all chosen by the caller. Field layout as the reference reads it:
frame header src/zflac.zig:343-374, subframe :426-433, LPC :499-520, residual :614-666, footer :546-548."""


class Bits:
    def __init__(self):
        self.acc, self.n = 0, 0

    def put(self, value: int, bits: int):
        assert 0 <= bits and -(1 << bits) < value < (1 << bits)
        self.acc = (self.acc << bits) | (value & ((1 << bits) - 1))
        self.n += bits
        return self

    def unary(self, zeros: int):
        return self.put(1, zeros + 1)

    def rice(self, value: int, k: int):
        """Signed residual, folded ((v << 1) ^ (v >> 63)), quotient in unary, k remainder bits (:656-663)."""
        cw = (value << 1) if value >= 0 else ((-value << 1) - 1)
        self.unary(cw >> k)
        return self.put(cw & ((1 << k) - 1), k) if k else self

    def align(self):
        return self.put(0, -self.n % 8)

    def bytes(self) -> bytes:
        assert self.n % 8 == 0
        return self.acc.to_bytes(self.n // 8, "big")


def crc8(data: bytes) -> int:
    c = 0
    for b in data:
        c ^= b
        for _ in range(8):
            c = ((c << 1) ^ 0x07) & 0xFF if c & 0x80 else (c << 1) & 0xFF
    return c


def crc16(data: bytes) -> int:
    c = 0
    for b in data:
        c ^= b << 8
        for _ in range(8):
            c = ((c << 1) ^ 0x8005) & 0xFFFF if c & 0x8000 else (c << 1) & 0xFFFF
    return c


def utf8_number(n: int) -> bytes:
    assert 0 <= n < (1 << 11)
    return bytes([n]) if n < 0x80 else bytes([0xC0 | (n >> 6), 0x80 | (n & 0x3F)])


DEPTH_CODE = {8: 1, 12: 2, 16: 4, 20: 5, 24: 6, 32: 7}


def lpc_mono_frame(number: int, depth: int, warmup, coefs, precision: int, shift: int, residuals, k: int = 0,
                   escape_bits: int = 0) -> bytes:
    """One fixed-block-size frame, one channel, one LPC subframe of order len(warmup) == len(coefs): block size =
    len(warmup) + len(residuals) (<= 256), 44.1 kHz, one partition: Rice with parameter k, or (escape_bits != 0) the
    escape code and every residual as a raw signed number of that many bits (:646-654)."""
    order = len(warmup)
    bs = order + len(residuals)
    assert 1 <= order <= 32 and len(coefs) == order and 1 <= bs <= 256 and 1 <= precision <= 15
    head = bytes([0xFF, 0xF8, (6 << 4) | 9, (0 << 4) | (DEPTH_CODE[depth] << 1)]) + utf8_number(number) + bytes([bs - 1])
    head += bytes([crc8(head)])
    b = Bits()
    b.put(0, 1).put(31 + order, 6).put(0, 1)  # subframe header: LPC of this order, no wasted bits
    for w in warmup:
        b.put(w, depth)
    b.put(precision - 1, 4).put(shift, 5)
    for c in coefs:
        b.put(c, precision)
    b.put(0, 2).put(0, 4)  # 4-bit Rice parameters, partition order 0
    if escape_bits:
        b.put(15, 4).put(escape_bits, 5)
        for r in residuals:
            b.put(r, escape_bits)
    else:
        b.put(k, 4)
        for r in residuals:
            b.rice(r, k)
    body = head + b.align().bytes()
    return body + crc16(body).to_bytes(2, "big")


def stream(frames, depth: int, channels: int, total: int, block_size: int, sample_rate: int = 44100, md5: bytes = bytes(16)) -> bytes:
    """fLaC + one last-block STREAMINFO + the frames. The reference compares the MD5 whatever it is (:267-280): decode once,
    take the digest the decoder reports with its InvalidChecksum, and build the stream again with it."""
    si = bytearray(34)
    si[0:2] = block_size.to_bytes(2, "big")
    si[2:4] = block_size.to_bytes(2, "big")
    packed = (sample_rate << 44) | ((channels - 1) << 41) | ((depth - 1) << 36) | total  # 20 + 3 + 5 + 36 bits
    si[10:18] = packed.to_bytes(8, "big")
    si[18:34] = md5
    return b"fLaC" + bytes([0x80, 0, 0, 34]) + bytes(si) + b"".join(frames)


def wide_sample_stream():
    """A 24-bit stream (32-bit container: the reference sums in i64, src/zflac.zig:314-319) whose LPC subframes reach samples
    of 30 bits through an escape-coded residual: 127 * s overflows int32, and the frame's depth + precision + log2(order)
    (24 + 8 + 0) says it cannot. Still inside i32, so the reference is defined — and keeps bits an int32 sum loses.
    40 frames of 3 samples, one channel; returned as a uint8 array with the MD5 the decoder computes."""
    import numpy as np
    from oracle import oracle

    def build(md5=bytes(16)):
        frames = [lpc_mono_frame(i, 24, [100 + i], [127], 8, 15, [(1 << 29) + 12345 + i, 7 - i], escape_bits=31) for i in range(40)]
        return np.frombuffer(stream(frames, 24, 1, 120, 3, md5=md5), dtype=np.uint8).copy()

    return build(oracle.decode(build()).md5_computed)
