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
    """The frame / sample number, coded like UTF-8 (:203-214)."""
    assert 0 <= n < (1 << 36)
    if n < 0x80:
        return bytes([n])
    for extra in range(1, 7):
        if n < (1 << (6 * extra + 6 - extra)):
            lead = (0xFF << (7 - extra)) & 0xFF
            return bytes([lead | (n >> (6 * extra))] + [0x80 | ((n >> (6 * k)) & 0x3F) for k in range(extra - 1, -1, -1)])
    raise AssertionError(n)


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


def stream(frames, depth: int, channels: int, total: int, block_size: int, sample_rate: int = 44100, md5: bytes = bytes(16),
           min_block_size=None) -> bytes:
    """fLaC + one last-block STREAMINFO + the frames. The reference compares the MD5 whatever it is (:267-280): decode once,
    take the digest the decoder reports with its InvalidChecksum, and build the stream again with it."""
    si = bytearray(34)
    si[0:2] = (min_block_size or block_size).to_bytes(2, "big")
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


# ---------------------------------------------------------------------------------------------------------------------
# A random encoder: any legal combination of the fields above, not what an audio encoder would choose. The PCM is drawn
# first and every subframe is coded to reproduce it exactly, so the expected output is known without any decoder.
# ---------------------------------------------------------------------------------------------------------------------
def _predict(x, coefs, shift, i):
    return sum(c * x[i - 1 - j] for j, c in enumerate(coefs)) >> shift  # Python ints: the reference's i64, floor shift


FIXED = {0: [], 1: [1], 2: [2, -1], 3: [3, -3, 1], 4: [4, -6, 4, -1]}


def _signal(rng, n, bits):
    """n values of `bits` signed bits: a random walk, a full-scale square wave, noise, silence, or full-scale noise with
    its low bits clear (wasted bits; on a side channel the shift back then leaves the container type: fuzz_case)."""
    import numpy as np
    lim = (1 << (bits - 1)) - 1
    kind = rng.integers(0, 6)
    if kind == 5:
        k = int(rng.integers(1, min(7, bits - 1)))
        return [int(t) for t in (rng.integers(-lim - 1, lim + 1, n) >> k) << k]
    if kind == 0:
        v = np.cumsum(rng.integers(-max(1, lim >> 6), max(1, lim >> 6) + 1, n))
    elif kind == 1:
        v = np.where((np.arange(n) // max(1, int(rng.integers(1, 9)))) % 2 == 0, lim, -lim - 1)
    elif kind == 2:
        v = rng.integers(-lim - 1, lim + 1, n)
    elif kind == 3:
        v = np.zeros(n, dtype=np.int64) + int(rng.integers(-lim - 1, lim + 1))
    else:
        v = rng.integers(-3, 4, n)
    return [int(t) for t in np.clip(v, -lim - 1, lim)]


def _put_residual(b: Bits, rng, res, order, bs):
    """Residual section (:614-666): coding method, partition order, per partition a Rice parameter or the escape code."""
    method = int(rng.integers(0, 2))
    if max([abs(r) for r in res], default=0).bit_length() > 26:
        method = 1  # 4-bit parameters stop at 14: unary runs of 2^13 zeros and more per residual make Python's big integers crawl
    pbits, kmax = (4, 14) if method == 0 else (5, 30)
    orders = [po for po in range(0, 9) if bs % (1 << po) == 0 and (bs >> po) >= order and (bs >> po) > 0]
    po = int(rng.choice(orders))
    b.put(method, 2).put(po, 4)
    psize, at = bs >> po, 0
    for part in range(1 << po):
        n = psize - order if part == 0 else psize
        rs = res[at:at + n]
        at += n
        mag = max([abs(r) for r in rs], default=0)
        need = 0 if mag == 0 and all(r == 0 for r in rs) else max(abs(r) if r >= 0 else abs(r) - 1 for r in rs).bit_length() + 1
        if rng.random() < 0.15 and need <= 31:
            b.put((1 << pbits) - 1, pbits).put(need, 5)
            for r in rs:
                if need:
                    b.put(r, need)
            continue
        ideal = max(0, mag.bit_length() - 1)
        k = int(min(kmax, max(0, ideal + int(rng.integers(-5, 3)))))
        k = max(k, ideal - 6)  # at most ~200 zeros in a unary run
        k = min(k, kmax)
        b.put(k, pbits)
        for r in rs:
            b.rice(r, k)


def _put_subframe(b: Bits, rng, x, depth, exact=False, bps=None):
    """One subframe reproducing x (values of `depth` bits) — or, unless `exact`, x with its low bits cleared (wasted bits)
    or held at its first value (CONSTANT); what was coded is returned. bps: the frame's bits per sample — what the
    reference reads a CONSTANT value with, on a side channel too (:447)."""
    bs = len(x)
    wasted = 0
    if depth > 4 and rng.random() < (0.6 if exact else 0.3):
        wasted = int(rng.integers(1, min(depth - 2, 9)))
        if exact:
            while wasted and any(v & ((1 << wasted) - 1) for v in x):
                wasted -= 1
        x = [(v >> wasted) << wasted for v in x]
    y = [v >> wasted for v in x]
    eff = depth - wasted
    kind = int(rng.choice(4, p=[0.1, 0.15, 0.3, 0.45]))
    ceff = (bps or depth) - wasted
    if kind == 0 and ((exact and any(v != y[0] for v in y)) or ceff < 1 or not -(1 << (ceff - 1)) <= y[0] < (1 << (ceff - 1))):
        kind = int(rng.integers(1, 4))
    if kind == 0:
        y = [y[0]] * bs
        x = [y[0] << wasted] * bs
    order = 0
    if kind == 2:
        order = int(rng.integers(0, min(4, bs) + 1))
    if kind == 3:
        order = int(rng.integers(1, min(32, bs) + 1))
    typ = {0: 0, 1: 1, 2: 8 + order, 3: 31 + order}[kind]
    b.put(0, 1).put(typ, 6).put(1 if wasted else 0, 1)
    if wasted:
        b.unary(wasted - 1)
    if kind == 0:
        b.put(y[0], ceff)
        return x
    if kind == 1:
        for v in y:
            b.put(v, eff)
        return x
    for v in y[:order]:
        b.put(v, eff)
    coefs, shift = FIXED.get(order, []), 0
    if kind == 3:
        precision = int(rng.integers(1, 16))
        shift = int(rng.integers(0, precision)) if rng.random() < 0.9 else int(rng.integers(0, 32))
        w = rng.normal(size=order)
        w = w / max(1e-9, abs(w).sum()) * rng.uniform(0.2, 1.6)
        lim = (1 << (precision - 1)) - 1
        coefs = [int(max(-lim - 1, min(lim, round(float(t) * (1 << shift))))) for t in w]
        b.put(precision - 1, 4).put(shift, 5)
        for c in coefs:
            b.put(c, precision)
    res = [y[i] - _predict(y, coefs, shift, i) for i in range(order, bs)]
    _put_residual(b, rng, res, order, bs)
    return x


def random_stream(seed: int, depths=(8, 12, 16, 16, 20, 24, 24), sizes=(16, 32, 33, 64, 100, 192, 255, 256, 576, 1024, 1152)):
    """(flac bytes as uint8 array, expected interleaved PCM before the MSB-align shift, channels, depth). depths: what the
    stream's depth is drawn from (the suite's seeds use the default; fuzz_hunt.py also runs 32)."""
    import hashlib
    import numpy as np
    rng = np.random.default_rng(seed)
    depth = int(rng.choice(list(depths)))
    ch_code = int(rng.choice([0, 1, 1, 2, 5, 7, 8, 9, 10, 10]))
    channels = 2 if ch_code >= 8 else ch_code + 1
    sizes = list(sizes)
    bs = int(rng.choice(sizes))
    n_frames = int(rng.integers(1, 6))
    last = int(rng.integers(1, bs + 1))
    # the sample rate as one of the table codes or spelled out after the frame number, in kHz, Hz or tens of Hz (:343-374)
    rate, rate_code, rate_tail = [(44100, 9, b""), (48000, 10, b""), (48000, 12, bytes([48])), (44100, 13, (44100).to_bytes(2, "big")),
                                  (48000, 14, (4800).to_bytes(2, "big")), (44100, 0, b"")][int(rng.integers(0, 6))]
    variable = rng.random() < 0.3  # blocking strategy bit: every frame its own size, numbered by its first sample (:343-374)
    depth_in_header = rng.random() < 0.8  # else code 0: "as STREAMINFO says" (:381-384)
    frames, pcm, used = [], [], []
    for f in range(n_frames):
        n = int(rng.choice(sizes)) if variable else last if (f == n_frames - 1 and n_frames > 1) else bs
        used.append(n)
        code = {192: 1, 576: 2, 1152: 3, 2304: 4, 4608: 5, 256: 8, 512: 9, 1024: 10, 2048: 11, 4096: 12, 8192: 13, 16384: 14,
                32768: 15}.get(n, 6 if n <= 256 else 7) if rng.random() < 0.7 else (6 if n <= 256 else 7)
        number = len(pcm) // channels if variable else f
        head = bytes([0xFF, 0xF9 if variable else 0xF8, (code << 4) | rate_code,
                      (ch_code << 4) | ((DEPTH_CODE[depth] if depth_in_header else 0) << 1)]) + utf8_number(number)
        head += bytes([n - 1]) if code == 6 else (n - 1).to_bytes(2, "big") if code == 7 else b""
        head += rate_tail
        head += bytes([crc8(head)])
        sig = [_signal(rng, n, depth) for _ in range(channels)]
        if ch_code >= 8 and rng.random() < 0.3:  # both channels full scale with the same low bits clear: a side channel with wasted bits
            k = int(rng.integers(1, min(7, depth - 1)))
            lim = (1 << (depth - 1)) - 1
            sig = [[int(t) for t in (rng.integers(-lim - 1, lim + 1, n) >> k) << k] for _ in range(2)]
        b = Bits()
        if ch_code < 8:
            out = [_put_subframe(b, rng, sig[c], depth) for c in range(channels)]
        else:
            L, R = sig
            side = [l - r for l, r in zip(L, R)]
            mid = [(l + r) >> 1 for l, r in zip(L, R)]
            plan = {8: ((L, depth), (side, depth + 1)), 9: ((side, depth + 1), (R, depth)), 10: ((mid, depth), (side, depth + 1))}[ch_code]
            a, s = (_put_subframe(b, rng, v, d, exact=True, bps=depth) for v, d in plan)
            if ch_code == 8:
                out = [a, [l - d for l, d in zip(a, s)]]
            elif ch_code == 9:
                out = [[d + r for d, r in zip(a, s)], s]
            else:
                m2 = [(m << 1) | (d & 1) for m, d in zip(a, s)]
                out = [[(m + d) >> 1 for m, d in zip(m2, s)], [(m - d) >> 1 for m, d in zip(m2, s)]]
        body = head + b.align().bytes()
        frames.append(body + crc16(body).to_bytes(2, "big"))
        for i in range(n):
            pcm.extend(out[c][i] for c in range(channels))
    nbytes = (depth + 7) // 8
    md5 = hashlib.md5(b"".join(int(v).to_bytes(nbytes, "little", signed=True) for v in pcm)).digest()
    total = len(pcm) // channels
    data = stream(frames, depth, channels, total, max(used), sample_rate=rate, md5=md5, min_block_size=min(used))
    return np.frombuffer(data, dtype=np.uint8).copy(), np.array(pcm, dtype=np.int64), channels, depth


def fuzz_case(seed, depths=None, sizes=None):
    """(stream, oracle result, exact) of random_stream(seed) — of random_stream(seed, depths) if given —, or None where the reference is undefined (the oracle's overflow
    flag: a side-channel value beyond the container, a sum beyond the reference's InterType). exact: the reference
    reproduces the PCM the stream was coded from; else one of its defined quirks changed it (a side channel's wasted-bits
    shift wraps in the container type, :494 / :537) — the stream's MD5 then does not match, and the oracle's samples
    (keep_on_md5_mismatch) are what a decoder without the gate must produce."""
    import numpy as np
    from oracle import oracle
    kw = dict(depths=depths) if depths else {}
    if sizes:
        kw["sizes"] = sizes
    data, pcm, channels, depth = random_stream(seed, **kw)
    ref = oracle.decode(data, trace_frames=True, keep_on_md5_mismatch=True)
    if ref.overflow_flag or ref.error == "UndefinedInReference":  # the latter: a predictor shift as wide as the InterType (:531)
        return None
    assert ref.error in ("OK", "InvalidChecksum"), (seed, ref.error)
    shift = 16 - depth if 9 <= depth <= 15 else 32 - depth if 17 <= depth <= 31 else 0
    exact = ref.error == "OK"
    assert not exact or np.array_equal(ref.samples.astype(np.int64), pcm << shift), seed
    return data, ref, exact
