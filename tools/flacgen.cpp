// flacgen.cpp — synthetic FLAC (RFC 9639) stream generator. TEST / BENCH INFRASTRUCTURE, not product code.
//
// BASELINE.json asks for test inputs "produced offline by a small in-repo Zig FLAC encoder that
// writes synthetic LPC/Rice-coded signals of the named shapes". There is no Zig toolchain in this
// image, so the encoder is C++ (same role, same outputs). It writes complete streams — fLaC
// marker, STREAMINFO with a correct MD5 and total, frames with valid CRC-8 / CRC-16 — for the
// shapes of BASELINE.json configs 2-5 and stays inside the behaviour the reference defines
// (SURVEY.md §7 Q8-Q10, Q16): no CONSTANT subframe on a side channel, side values that fit the
// container, predictor sums that fit the reference's accumulator type.
//
// Everything is a pure function of fg_params (seeded xoshiro256**, per-frame streams), so the
// same parameters give the same bytes on any machine and with any thread count.
//
// Build: g++ -O3 -std=c++17 -shared -fPIC -pthread (tools/flacgen.py does it).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

extern "C" {

typedef struct fg_params {
    uint32_t struct_size;
    uint32_t sample_rate;
    uint32_t channels;             // 1..8
    uint32_t bps;                  // 4..32 (STREAMINFO and frame headers agree)
    uint32_t block_size;           // fixed block size (variable_blocksize = 0)
    uint32_t lpc_order;            // 1..32: order of LPC subframes
    uint32_t qlp_precision;        // 1..15
    uint32_t rice_method;          // 0: 4-bit parameters, 1: 5-bit parameters
    uint32_t min_partition_order;
    uint32_t max_partition_order;  // the encoder picks the cheapest valid order in [min, max]
    uint32_t stereo_mode;          // 0 independent, 1 left/side, 2 side/right, 3 mid/side, 4 cycle 0..3 per frame
    uint32_t subframe_mix;         // 0: every subframe LPC(lpc_order); 1: LPC / FIXED 0-4 / VERBATIM / CONSTANT mix
    uint32_t wasted_bits;          // 0 none; 1: frames carry 1..9 wasted bits (cycling)
    uint32_t variable_blocksize;   // 0 fixed; 1: blocking strategy 1, sizes cycle 16/192/576/1000/4096/4608/16384
    uint32_t escape_every;         // 0 never; n: every n-th partition is written escaped (raw)
    uint32_t signal;               // 0: tones + shaped noise; 1: wide dynamic range (noise level sweeps with the frame)
    uint32_t set_total;            // 1: STREAMINFO total samples set; 0: unknown (0)
    uint32_t n_threads;            // 0 = hardware concurrency
    uint64_t n_samples;            // per channel
    uint64_t seed;
    double amplitude;              // tone amplitude as a fraction of full scale (0..0.45)
    double noise;                  // noise sigma as a fraction of full scale
} fg_params;

typedef struct fg_result {
    uint8_t *flac;
    uint64_t flac_bytes;
    void *pcm;            // interleaved, in the container the reference returns (i8/i16/i32), MSB-aligned (Q2)
    uint64_t pcm_bytes;
    uint64_t n_frames;
    uint64_t audio_offset;
    uint32_t container_bytes;
    uint32_t pad;
    uint8_t md5[16];
} fg_result;

int fg_generate(const fg_params *p, int want_pcm, fg_result *out);
int fg_generate_batch(const fg_params *ps, uint32_t n, int want_pcm, fg_result *outs, uint32_t n_threads);
void fg_free(fg_result *r);
}

namespace {

// ------------------------------------------------------------------ RNG
struct Xo {
    uint64_t s[4];
    static uint64_t splitmix(uint64_t &x) {
        uint64_t z = (x += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    explicit Xo(uint64_t seed) {
        for (auto &v : s) v = splitmix(seed);
    }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next() {
        const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return r;
    }
    // ~N(0,1): sum of four 16-bit uniforms, centred and scaled (Irwin-Hall, var = 4/12)
    double gauss() {
        const uint64_t r = next();
        const double u = (double)(r & 0xFFFF) + (double)((r >> 16) & 0xFFFF) + (double)((r >> 32) & 0xFFFF) + (double)(r >> 48);
        return (u - 2.0 * 65535.0) * (1.0 / (65536.0 * 0.5773502691896258));
    }
};

// ------------------------------------------------------------------ checksums
uint8_t g_crc8[256];
uint16_t g_crc16[8][256];
struct TableInit {
    TableInit() {
        for (int i = 0; i < 256; ++i) {
            uint32_t c = (uint32_t)i;
            for (int k = 0; k < 8; ++k) c = (c & 0x80) ? ((c << 1) ^ 0x07) & 0xFF : (c << 1) & 0xFF;
            g_crc8[i] = (uint8_t)c;
            uint32_t d = (uint32_t)i << 8;
            for (int k = 0; k < 8; ++k) d = (d & 0x8000) ? ((d << 1) ^ 0x8005) & 0xFFFF : (d << 1) & 0xFFFF;
            g_crc16[0][i] = (uint16_t)d;
        }
        for (int t = 1; t < 8; ++t)
            for (int i = 0; i < 256; ++i) {
                const uint16_t v = g_crc16[t - 1][i];
                g_crc16[t][i] = (uint16_t)((v << 8) ^ g_crc16[0][v >> 8]);
            }
    }
} g_table_init;

uint8_t crc8(const uint8_t *p, size_t n) {
    uint8_t c = 0;
    for (size_t i = 0; i < n; ++i) c = g_crc8[c ^ p[i]];
    return c;
}
uint16_t crc16(const uint8_t *p, size_t n) {
    uint16_t c = 0;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {  // slice-by-8
        const uint8_t b0 = p[i] ^ (uint8_t)(c >> 8), b1 = p[i + 1] ^ (uint8_t)c;
        c = g_crc16[7][b0] ^ g_crc16[6][b1] ^ g_crc16[5][p[i + 2]] ^ g_crc16[4][p[i + 3]] ^ g_crc16[3][p[i + 4]] ^
            g_crc16[2][p[i + 5]] ^ g_crc16[1][p[i + 6]] ^ g_crc16[0][p[i + 7]];
    }
    for (; i < n; ++i) c = (uint16_t)((c << 8) ^ g_crc16[0][(c >> 8) ^ p[i]]);
    return c;
}

struct Md5 {
    uint32_t a = 0x67452301u, b = 0xefcdab89u, c = 0x98badcfeu, d = 0x10325476u;
    uint64_t total = 0;
    uint8_t buf[64];
    size_t fill = 0;
    static uint32_t rl(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }
    void block(const uint8_t *p) {
        uint32_t m[16];
        memcpy(m, p, 64);
        uint32_t A = a, B = b, C = c, D = d;
#define FG_F(x, y, z) ((z) ^ ((x) & ((y) ^ (z))))
#define FG_G(x, y, z) ((y) ^ ((z) & ((x) ^ (y))))
#define FG_H(x, y, z) ((x) ^ (y) ^ (z))
#define FG_I(x, y, z) ((y) ^ ((x) | ~(z)))
#define FG_S(f, w, x, y, z, g, k, s) w = x + rl(w + f(x, y, z) + m[g] + k, s)
        FG_S(FG_F, A, B, C, D, 0, 0xd76aa478u, 7);  FG_S(FG_F, D, A, B, C, 1, 0xe8c7b756u, 12);
        FG_S(FG_F, C, D, A, B, 2, 0x242070dbu, 17); FG_S(FG_F, B, C, D, A, 3, 0xc1bdceeeu, 22);
        FG_S(FG_F, A, B, C, D, 4, 0xf57c0fafu, 7);  FG_S(FG_F, D, A, B, C, 5, 0x4787c62au, 12);
        FG_S(FG_F, C, D, A, B, 6, 0xa8304613u, 17); FG_S(FG_F, B, C, D, A, 7, 0xfd469501u, 22);
        FG_S(FG_F, A, B, C, D, 8, 0x698098d8u, 7);  FG_S(FG_F, D, A, B, C, 9, 0x8b44f7afu, 12);
        FG_S(FG_F, C, D, A, B, 10, 0xffff5bb1u, 17); FG_S(FG_F, B, C, D, A, 11, 0x895cd7beu, 22);
        FG_S(FG_F, A, B, C, D, 12, 0x6b901122u, 7); FG_S(FG_F, D, A, B, C, 13, 0xfd987193u, 12);
        FG_S(FG_F, C, D, A, B, 14, 0xa679438eu, 17); FG_S(FG_F, B, C, D, A, 15, 0x49b40821u, 22);
        FG_S(FG_G, A, B, C, D, 1, 0xf61e2562u, 5);  FG_S(FG_G, D, A, B, C, 6, 0xc040b340u, 9);
        FG_S(FG_G, C, D, A, B, 11, 0x265e5a51u, 14); FG_S(FG_G, B, C, D, A, 0, 0xe9b6c7aau, 20);
        FG_S(FG_G, A, B, C, D, 5, 0xd62f105du, 5);  FG_S(FG_G, D, A, B, C, 10, 0x02441453u, 9);
        FG_S(FG_G, C, D, A, B, 15, 0xd8a1e681u, 14); FG_S(FG_G, B, C, D, A, 4, 0xe7d3fbc8u, 20);
        FG_S(FG_G, A, B, C, D, 9, 0x21e1cde6u, 5);  FG_S(FG_G, D, A, B, C, 14, 0xc33707d6u, 9);
        FG_S(FG_G, C, D, A, B, 3, 0xf4d50d87u, 14); FG_S(FG_G, B, C, D, A, 8, 0x455a14edu, 20);
        FG_S(FG_G, A, B, C, D, 13, 0xa9e3e905u, 5); FG_S(FG_G, D, A, B, C, 2, 0xfcefa3f8u, 9);
        FG_S(FG_G, C, D, A, B, 7, 0x676f02d9u, 14); FG_S(FG_G, B, C, D, A, 12, 0x8d2a4c8au, 20);
        FG_S(FG_H, A, B, C, D, 5, 0xfffa3942u, 4);  FG_S(FG_H, D, A, B, C, 8, 0x8771f681u, 11);
        FG_S(FG_H, C, D, A, B, 11, 0x6d9d6122u, 16); FG_S(FG_H, B, C, D, A, 14, 0xfde5380cu, 23);
        FG_S(FG_H, A, B, C, D, 1, 0xa4beea44u, 4);  FG_S(FG_H, D, A, B, C, 4, 0x4bdecfa9u, 11);
        FG_S(FG_H, C, D, A, B, 7, 0xf6bb4b60u, 16); FG_S(FG_H, B, C, D, A, 10, 0xbebfbc70u, 23);
        FG_S(FG_H, A, B, C, D, 13, 0x289b7ec6u, 4); FG_S(FG_H, D, A, B, C, 0, 0xeaa127fau, 11);
        FG_S(FG_H, C, D, A, B, 3, 0xd4ef3085u, 16); FG_S(FG_H, B, C, D, A, 6, 0x04881d05u, 23);
        FG_S(FG_H, A, B, C, D, 9, 0xd9d4d039u, 4);  FG_S(FG_H, D, A, B, C, 12, 0xe6db99e5u, 11);
        FG_S(FG_H, C, D, A, B, 15, 0x1fa27cf8u, 16); FG_S(FG_H, B, C, D, A, 2, 0xc4ac5665u, 23);
        FG_S(FG_I, A, B, C, D, 0, 0xf4292244u, 6);  FG_S(FG_I, D, A, B, C, 7, 0x432aff97u, 10);
        FG_S(FG_I, C, D, A, B, 14, 0xab9423a7u, 15); FG_S(FG_I, B, C, D, A, 5, 0xfc93a039u, 21);
        FG_S(FG_I, A, B, C, D, 12, 0x655b59c3u, 6); FG_S(FG_I, D, A, B, C, 3, 0x8f0ccc92u, 10);
        FG_S(FG_I, C, D, A, B, 10, 0xffeff47du, 15); FG_S(FG_I, B, C, D, A, 1, 0x85845dd1u, 21);
        FG_S(FG_I, A, B, C, D, 8, 0x6fa87e4fu, 6);  FG_S(FG_I, D, A, B, C, 15, 0xfe2ce6e0u, 10);
        FG_S(FG_I, C, D, A, B, 6, 0xa3014314u, 15); FG_S(FG_I, B, C, D, A, 13, 0x4e0811a1u, 21);
        FG_S(FG_I, A, B, C, D, 4, 0xf7537e82u, 6);  FG_S(FG_I, D, A, B, C, 11, 0xbd3af235u, 10);
        FG_S(FG_I, C, D, A, B, 2, 0x2ad7d2bbu, 15); FG_S(FG_I, B, C, D, A, 9, 0xeb86d391u, 21);
#undef FG_S
        a += A; b += B; c += C; d += D;
    }
    void update(const uint8_t *p, size_t n) {
        total += n;
        if (fill) {
            const size_t k = std::min(n, 64 - fill);
            memcpy(buf + fill, p, k);
            fill += k; p += k; n -= k;
            if (fill == 64) { block(buf); fill = 0; }
        }
        for (; n >= 64; p += 64, n -= 64) block(p);
        if (n) { memcpy(buf, p, n); fill = n; }
    }
    void finish(uint8_t out[16]) {
        const uint64_t bits = total * 8;
        uint8_t pad[72] = {0x80};
        const size_t padn = (fill < 56) ? 56 - fill : 120 - fill;
        update(pad, padn);
        uint8_t lenb[8];
        for (int i = 0; i < 8; ++i) lenb[i] = (uint8_t)(bits >> (8 * i));
        update(lenb, 8);
        const uint32_t v[4] = {a, b, c, d};
        for (int i = 0; i < 4; ++i)
            for (int k = 0; k < 4; ++k) out[4 * i + k] = (uint8_t)(v[i] >> (8 * k));
    }
};

// ------------------------------------------------------------------ bit writer (MSB first)
struct BitWriter {
    std::vector<uint8_t> &out;
    uint64_t acc = 0;
    int n = 0;  // bits held in acc (low n bits)
    explicit BitWriter(std::vector<uint8_t> &o) : out(o) {}
    void put(uint64_t v, int bits) {  // bits in [0, 32]
        if (bits == 0) return;
        acc = (acc << bits) | (v & ((bits == 64) ? ~0ull : ((1ull << bits) - 1)));
        n += bits;
        while (n >= 8) {
            out.push_back((uint8_t)(acc >> (n - 8)));
            n -= 8;
        }
    }
    void put_signed(int64_t v, int bits) {
        if (bits > 32) {
            put((uint64_t)v >> 32, bits - 32);
            put((uint64_t)v & 0xFFFFFFFFull, 32);
        } else put((uint64_t)v, bits);
    }
    void put_unary(uint32_t q) {  // q zeros then a one
        while (q >= 32) { put(0, 32); q -= 32; }
        put(1, (int)q + 1);
    }
    void align() {
        if (n) put(0, 8 - n);
    }
};

int bits_needed_signed(int64_t lo, int64_t hi) {  // two's complement width holding [lo, hi]
    int b = 1;
    while (b < 64 && (lo < -(1ll << (b - 1)) || hi > (1ll << (b - 1)) - 1)) ++b;
    return b;
}

// ------------------------------------------------------------------ frame plan
struct FramePlan {
    uint64_t first_sample;
    uint32_t block_size;
    uint64_t index;  // frame number
};

const uint32_t kVarSizes[7] = {16, 192, 576, 1000, 4096, 4608, 16384};

std::vector<FramePlan> plan_frames(const fg_params &p) {
    std::vector<FramePlan> v;
    uint64_t pos = 0, idx = 0;
    while (pos < p.n_samples) {
        uint32_t bs = p.variable_blocksize ? kVarSizes[idx % 7] : p.block_size;
        if ((uint64_t)bs > p.n_samples - pos) bs = (uint32_t)(p.n_samples - pos);
        v.push_back({pos, bs, idx});
        pos += bs;
        ++idx;
    }
    // the reference accepts a block size of 1 only on the last frame (zflac.zig:405): true by construction
    return v;
}

// ------------------------------------------------------------------ signal
// Samples of one frame, all channels, as exact integers of `bps` bits.
void synth_frame(const fg_params &p, const FramePlan &f, std::vector<std::vector<int64_t>> &ch) {
    const uint32_t C = p.channels, n = f.block_size;
    const double full = std::ldexp(1.0, (int)p.bps - 1);
    Xo rng(p.seed * 0x9E3779B97F4A7C15ull + f.index * 0xD1B54A32D192ED03ull + 1);
    double amp = p.amplitude * full, sigma = p.noise * full;
    if (p.signal == 1) {  // wide dynamic range: noise level sweeps over the frames, tones stay
        const int span = std::max(1, (int)p.bps - 4);
        sigma = std::ldexp(0.6, (int)(f.index % (uint64_t)span)) ;
        if (sigma > 0.2 * full) sigma = 0.2 * full;
    }
    const double limit = (C == 2 && p.stereo_mode != 0) ? 0.49 * full : 0.98 * full;  // side must fit the container (Q9)
    static const double kFreq[3] = {220.0, 1318.5, 5274.0};
    static const double kRel[3] = {1.0, 0.5, 0.25};
    ch.assign(C, std::vector<int64_t>(n));
    std::vector<double> base(n);
    for (uint32_t c = 0; c < C; ++c) {
        const bool follow = (C == 2 && c == 1);  // R = 0.8 L + its own noise: gives mid/side something to win
        // tones by phasor rotation from the absolute sample index (continuous across frames)
        double re[3], im[3], cr[3], ci[3];
        for (int t = 0; t < 3; ++t) {
            const double w = 2.0 * M_PI * kFreq[t] * (1.0 + 0.013 * c) / (double)p.sample_rate;
            const double ph = std::fmod(w * (double)f.first_sample, 2.0 * M_PI) + 0.7 * c;
            re[t] = std::cos(ph); im[t] = std::sin(ph);
            cr[t] = std::cos(w); ci[t] = std::sin(w);
        }
        double z1 = 0, z2 = 0;  // AR(2) noise shaping: y = e + 1.2 z1 - 0.5 z2  (spectral tilt)
        const double ar_gain = 0.45;
        const double sg = follow ? sigma * 0.66 : sigma;
        for (uint32_t i = 0; i < n; ++i) {
            double v = 0;
            if (!follow) {
                for (int t = 0; t < 3; ++t) {
                    v += kRel[t] * im[t];
                    const double nr = re[t] * cr[t] - im[t] * ci[t];
                    im[t] = re[t] * ci[t] + im[t] * cr[t];
                    re[t] = nr;
                }
                v *= amp / 1.75;
            } else {
                v = 0.8 * base[i];
            }
            const double e = rng.gauss();
            const double y = e + 1.2 * z1 - 0.5 * z2;
            z2 = z1; z1 = y;
            v += sg * ar_gain * y;
            if (c == 0) base[i] = v;
            if (v > limit - 1) v = limit - 1;
            if (v < -limit) v = -limit;
            ch[c][i] = (int64_t)std::llrint(v);
        }
    }
    if (p.subframe_mix) {  // some frames carry a constant channel (never one that becomes a side channel, Q8)
        if (f.index % 5 == 2) {
            const uint32_t c = (uint32_t)((f.index / 5) % C);
            const int64_t v = ch[c][0];
            std::fill(ch[c].begin(), ch[c].end(), v);
        }
    }
    if (p.wasted_bits) {  // low bits cleared on every channel of the frame
        const int w = 1 + (int)(f.index % 9);
        if (w < (int)p.bps - 2)
            for (auto &s : ch)
                for (auto &x : s) x = (x >> w) * ((int64_t)1 << w);
    }
}

// ------------------------------------------------------------------ subframe encoding
struct RicePlan {
    uint32_t order;               // partition order
    std::vector<uint32_t> param;  // per partition; 0xFFFFFFFF = escaped
    std::vector<uint32_t> esc_bits;
    uint64_t bits;
};

inline uint64_t zigzag(int64_t r) { return ((uint64_t)r << 1) ^ (uint64_t)(r >> 63); }

// Cheapest partitioning of the residual res[0..n) (first `pred_order` entries absent).
RicePlan plan_rice(const fg_params &p, const std::vector<int64_t> &res, uint32_t bs, uint32_t pred_order, uint64_t salt) {
    const uint32_t kmax = p.rice_method ? 30 : 14;
    const uint32_t pbits = p.rice_method ? 5 : 4;
    RicePlan best;
    best.bits = ~0ull;
    auto valid = [&](uint32_t po) {  // must divide the block; first partition may not be negative (Q16)
        return ((bs >> po) << po) == bs && (bs >> po) >= pred_order && (bs >> po) > 0;
    };
    uint32_t lo = p.min_partition_order, hi = std::min<uint32_t>(p.max_partition_order, 15);
    while (lo > 0 && !valid(lo)) --lo;  // short last block: fall back to the largest valid order below
    for (uint32_t po = lo; po <= hi; ++po) {
        if (!valid(po)) break;
        RicePlan rp;
        rp.order = po;
        rp.bits = 0;
        const uint32_t np = 1u << po, psz = bs >> po;
        rp.param.resize(np);
        rp.esc_bits.assign(np, 0);
        for (uint32_t q = 0; q < np; ++q) {
            const uint32_t a = (q == 0) ? pred_order : q * psz, b = (q + 1) * psz;
            const uint32_t cnt = b - a;
            uint64_t sum = 0;
            int64_t mn = 0, mx = 0;
            for (uint32_t i = a; i < b; ++i) {
                sum += zigzag(res[i]);
                mn = std::min(mn, res[i]);
                mx = std::max(mx, res[i]);
            }
            uint32_t k = 0;
            if (cnt) {
                const uint64_t mean = sum / cnt;
                while (k < 40 && (mean >> (k + 1)) != 0) ++k;  // k ~ floor(log2(mean)); k+1 is tried below
            }
            const bool force_esc = p.escape_every && ((salt + q) % p.escape_every == p.escape_every - 1);
            if (k > kmax || force_esc) {
                rp.param[q] = 0xFFFFFFFFu;
                const int eb = (mn == 0 && mx == 0) ? 0 : bits_needed_signed(mn, mx);
                rp.esc_bits[q] = (uint32_t)eb;
                rp.bits += pbits + 5 + (uint64_t)eb * cnt;
            } else {
                // exact cost for k and k+1, keep the cheaper
                uint64_t bk[2] = {0, 0};
                for (int t = 0; t < 2; ++t) {
                    const uint32_t kk = std::min(kmax, k + (uint32_t)t);
                    uint64_t s = 0;
                    for (uint32_t i = a; i < b; ++i) s += (zigzag(res[i]) >> kk) + 1 + kk;
                    bk[t] = s;
                }
                if (bk[1] < bk[0] && k + 1 <= kmax) { k += 1; rp.bits += bk[1]; } else rp.bits += bk[0];
                rp.param[q] = k;
                rp.bits += pbits;
            }
        }
        if (rp.bits < best.bits) best = std::move(rp);
    }
    return best;
}

void write_residual(const fg_params &p, BitWriter &bw, const RicePlan &rp, const std::vector<int64_t> &res, uint32_t bs,
                    uint32_t pred_order) {
    const uint32_t pbits = p.rice_method ? 5 : 4;
    bw.put(p.rice_method ? 1 : 0, 2);
    bw.put(rp.order, 4);
    const uint32_t np = 1u << rp.order, psz = bs >> rp.order;
    for (uint32_t q = 0; q < np; ++q) {
        const uint32_t a = (q == 0) ? pred_order : q * psz, b = (q + 1) * psz;
        if (rp.param[q] == 0xFFFFFFFFu) {
            bw.put((1u << pbits) - 1, (int)pbits);
            bw.put(rp.esc_bits[q], 5);
            if (rp.esc_bits[q])
                for (uint32_t i = a; i < b; ++i) bw.put_signed(res[i], (int)rp.esc_bits[q]);
        } else {
            const uint32_t k = rp.param[q];
            bw.put(k, (int)pbits);
            for (uint32_t i = a; i < b; ++i) {
                const uint64_t u = zigzag(res[i]);
                const uint64_t qv = u >> k;
                if (qv + 1 + k <= 32) {
                    bw.put((1ull << k) | (u & ((1ull << k) - 1)), (int)(qv + 1 + k));
                } else {
                    bw.put_unary((uint32_t)qv);
                    bw.put(u & ((1ull << k) - 1), (int)k);
                }
            }
        }
    }
}

// Quantised LPC coefficients for s[0..n) (libFLAC-style: windowed autocorrelation, Levinson-Durbin,
// shift chosen so the largest coefficient uses the full precision, error feedback on rounding).
bool fit_lpc(const std::vector<int64_t> &s, uint32_t n, uint32_t order, uint32_t precision, int32_t *qcoef, uint32_t &shift) {
    std::vector<double> w(n);
    for (uint32_t i = 0; i < n; ++i) {
        const double t = (double)i / (double)(n - 1 ? n - 1 : 1);
        const double win = 0.5 - 0.5 * std::cos(2.0 * M_PI * t);  // Hann
        w[i] = win * (double)s[i];
    }
    double ac[33];
    for (uint32_t lag = 0; lag <= order; ++lag) {
        double a = 0;
        for (uint32_t i = lag; i < n; ++i) a += w[i] * w[i - lag];
        ac[lag] = a;
    }
    if (ac[0] <= 0) return false;
    ac[0] *= 1.0 + 1e-9;
    double lpc[33] = {0}, tmp[33];
    double err = ac[0];
    for (uint32_t i = 0; i < order; ++i) {
        double r = -ac[i + 1];
        for (uint32_t j = 0; j < i; ++j) r -= lpc[j] * ac[i - j];
        r /= err;
        memcpy(tmp, lpc, sizeof(double) * i);
        lpc[i] = r;
        for (uint32_t j = 0; j < i; ++j) lpc[j] = tmp[j] + r * tmp[i - 1 - j];
        err *= (1.0 - r * r);
        if (err <= 0) { err = 1e-9; }
    }
    double cmax = 0;
    for (uint32_t i = 0; i < order; ++i) cmax = std::max(cmax, std::fabs(lpc[i]));
    if (cmax <= 0) return false;
    int log2c;
    std::frexp(cmax, &log2c);  // cmax = m * 2^log2c, m in [0.5, 1)
    int sh = (int)precision - log2c - 1;
    if (sh > 15) sh = 15;
    if (sh < 0) return false;  // would need a negative shift: not representable for the reference (Q17)
    const int32_t qmax = (1 << (precision - 1)) - 1, qmin = -(1 << (precision - 1));
    double errf = 0;
    for (uint32_t i = 0; i < order; ++i) {
        errf += -lpc[i] * (double)(1 << sh);
        long q = std::lrint(errf);
        if (q > qmax) q = qmax;
        if (q < qmin) q = qmin;
        qcoef[i] = (int32_t)q;
        errf -= (double)q;
    }
    shift = (uint32_t)sh;
    return true;
}

const int kFixed[5][4] = {{0, 0, 0, 0}, {1, 0, 0, 0}, {2, -1, 0, 0}, {3, -3, 1, 0}, {4, -6, 4, -1}};

// Encodes one subframe of `depth` bits. `allow_constant` is false on side channels (Q8).
// acc_bits: width of the reference's accumulator type (16/32/64, Q10): predictors whose worst-case sum
// does not fit are not used.
void encode_subframe(const fg_params &p, BitWriter &bw, const std::vector<int64_t> &sig, uint32_t bs, uint32_t depth,
                     bool allow_constant, uint32_t acc_bits, uint64_t salt) {
    // wasted bits
    uint32_t wasted = 0;
    {
        uint64_t orv = 0;
        for (uint32_t i = 0; i < bs; ++i) orv |= (uint64_t)sig[i];
        if (orv != 0) {
            while (((orv >> wasted) & 1) == 0) ++wasted;
        }
        if (wasted >= depth) wasted = 0;
    }
    std::vector<int64_t> s(bs);
    for (uint32_t i = 0; i < bs; ++i) s[i] = sig[i] >> wasted;
    const uint32_t eff = depth - wasted;
    bool constant = true;
    for (uint32_t i = 1; i < bs && constant; ++i) constant = s[i] == s[0];

    auto header = [&](uint32_t type) {
        bw.put(0, 1);
        bw.put(type, 6);
        bw.put(wasted ? 1 : 0, 1);
        if (wasted) bw.put_unary(wasted - 1);
    };
    auto verbatim = [&]() {
        header(1);
        for (uint32_t i = 0; i < bs; ++i) bw.put_signed(s[i], (int)eff);
    };

    if (constant && allow_constant) {
        // the reference reads bps - wasted bits here, which equals eff only off the side channel (Q8)
        header(0);
        bw.put_signed(s[0], (int)eff);
        return;
    }
    // choose the predictor
    uint32_t kind = 2, order = p.lpc_order;  // 0 verbatim, 1 fixed, 2 lpc
    if (p.subframe_mix) {
        const uint64_t sel = salt % 8;
        if (sel == 7) kind = 0;
        else if (sel >= 2 && sel <= 6) { kind = 1; order = (uint32_t)(sel - 2); }
        else if (sel == 1) order = 1 + (uint32_t)((salt / 8) % p.lpc_order);
    }
    if (order > 32) order = 32;
    if (kind == 2 && (order == 0 || bs <= order)) kind = 0;
    if (kind == 1 && bs <= order) kind = 0;
    if (kind == 0) { verbatim(); return; }

    int32_t qc[32] = {0};
    uint32_t shift = 0;
    if (kind == 2) {
        uint32_t prec = p.qlp_precision;
        // keep |sum| <= 2^(eff-1 + prec-1 + ceil(log2(order))) inside the reference's accumulator (Q10)
        uint32_t lg = 0;
        while ((1u << lg) < order) ++lg;
        while (prec > 1 && eff + prec + lg > acc_bits) --prec;
        if (eff + prec + lg > acc_bits || !fit_lpc(s, bs, order, prec, qc, shift)) { kind = 1; order = std::min<uint32_t>(2, order); }
        else {
            // store the precision actually used
            std::vector<int64_t> res(bs, 0);
            bool ok = true;
            for (uint32_t i = order; i < bs && ok; ++i) {
                int64_t acc = 0;
                for (uint32_t j = 0; j < order; ++j) acc += (int64_t)qc[j] * s[i - 1 - j];
                const int64_t r = s[i] - (acc >> shift);
                if (r > 0x3FFFFFFFll || r < -0x40000000ll) ok = false;
                res[i] = r;
            }
            if (!ok) { verbatim(); return; }
            header(32 + order - 1);
            for (uint32_t i = 0; i < order; ++i) bw.put_signed(s[i], (int)eff);
            bw.put(prec - 1, 4);
            bw.put(shift, 5);
            for (uint32_t j = 0; j < order; ++j) bw.put_signed(qc[j], (int)prec);
            const RicePlan rp = plan_rice(p, res, bs, order, salt);
            write_residual(p, bw, rp, res, bs, order);
            return;
        }
    }
    // fixed predictor
    if (eff + 4 > acc_bits - 1) { verbatim(); return; }
    std::vector<int64_t> res(bs, 0);
    bool ok = true;
    for (uint32_t i = order; i < bs && ok; ++i) {
        int64_t acc = 0;
        for (uint32_t j = 0; j < order; ++j) acc += (int64_t)kFixed[order][j] * s[i - 1 - j];
        const int64_t r = s[i] - acc;
        if (r > 0x3FFFFFFFll || r < -0x40000000ll) ok = false;
        res[i] = r;
    }
    if (!ok) { verbatim(); return; }
    header(8 + order);
    for (uint32_t i = 0; i < order; ++i) bw.put_signed(s[i], (int)eff);
    const RicePlan rp = plan_rice(p, res, bs, order, salt);
    write_residual(p, bw, rp, res, bs, order);
}

void put_utf8(std::vector<uint8_t> &o, uint64_t v) {
    if (v < 0x80) { o.push_back((uint8_t)v); return; }
    int n = 2;
    while (n < 7 && v >= (1ull << (5 * n + 1))) ++n;  // n bytes carry 5n+1 bits
    o.push_back((uint8_t)((0xFF << (8 - n)) | (v >> (6 * (n - 1)))));
    for (int i = n - 2; i >= 0; --i) o.push_back((uint8_t)(0x80 | ((v >> (6 * i)) & 0x3F)));
}

uint32_t container_of(uint32_t bps) { return bps <= 8 ? 1 : bps <= 16 ? 2 : 4; }

// One whole frame -> bytes appended to `out`; PCM (container-typed, MSB aligned) -> pcm if non-null.
void encode_frame(const fg_params &p, const FramePlan &f, std::vector<uint8_t> &out, uint8_t *pcm_base,
                  std::vector<std::vector<int64_t>> &ch) {
    synth_frame(p, f, ch);
    const uint32_t C = p.channels, bs = f.block_size;
    const uint32_t cb = container_of(p.bps);
    if (pcm_base) {
        const uint32_t msb = (p.bps >= 9 && p.bps <= 15) ? 16 - p.bps : (p.bps >= 17 && p.bps <= 31) ? 32 - p.bps : 0;
        uint8_t *dst = pcm_base + f.first_sample * C * cb;
        for (uint32_t i = 0; i < bs; ++i)
            for (uint32_t c = 0; c < C; ++c) {
                const int64_t v = ch[c][i] * ((int64_t)1 << msb);
                if (cb == 1) ((int8_t *)dst)[(size_t)i * C + c] = (int8_t)v;
                else if (cb == 2) ((int16_t *)dst)[(size_t)i * C + c] = (int16_t)v;
                else ((int32_t *)dst)[(size_t)i * C + c] = (int32_t)v;
            }
    }
    // channel assignment
    uint32_t mode = 0;
    if (C == 2) {
        mode = p.stereo_mode == 4 ? (uint32_t)(f.index % 4) : p.stereo_mode;
        if (p.bps == 32) mode = 0;  // a 33-bit side channel does not fit the reference's container (Q9)
    }
    const uint32_t ch_code = (C == 2 && mode) ? 7 + mode : C - 1;
    std::vector<int64_t> a, b;
    uint32_t side_channel = 99;
    if (mode == 1) {  // left/side
        b.resize(bs);
        for (uint32_t i = 0; i < bs; ++i) b[i] = ch[0][i] - ch[1][i];
        side_channel = 1;
    } else if (mode == 2) {  // side/right
        a.resize(bs);
        for (uint32_t i = 0; i < bs; ++i) a[i] = ch[0][i] - ch[1][i];
        side_channel = 0;
    } else if (mode == 3) {  // mid/side
        a.resize(bs); b.resize(bs);
        for (uint32_t i = 0; i < bs; ++i) {
            a[i] = (ch[0][i] + ch[1][i]) >> 1;
            b[i] = ch[0][i] - ch[1][i];
        }
        side_channel = 1;
    }
    const size_t start = out.size();
    // header
    out.push_back(0xFF);
    out.push_back(p.variable_blocksize ? 0xF9 : 0xF8);
    uint32_t bs_code, bs_extra = 0;
    if (bs == 192) bs_code = 1;
    else if (bs == 576 || bs == 1152 || bs == 2304 || bs == 4608) bs_code = bs == 576 ? 2 : bs == 1152 ? 3 : bs == 2304 ? 4 : 5;
    else if (bs >= 256 && bs <= 32768 && (bs & (bs - 1)) == 0) {
        bs_code = 0;
        while ((1u << bs_code) != bs) ++bs_code;
    } else if (bs <= 256) { bs_code = 6; bs_extra = 1; }
    else { bs_code = 7; bs_extra = 2; }
    uint32_t sr_code = 0;
    switch (p.sample_rate) {
        case 88200: sr_code = 1; break; case 176400: sr_code = 2; break; case 192000: sr_code = 3; break;
        case 8000: sr_code = 4; break; case 16000: sr_code = 5; break; case 22050: sr_code = 6; break;
        case 24000: sr_code = 7; break; case 32000: sr_code = 8; break; case 44100: sr_code = 9; break;
        case 48000: sr_code = 10; break; case 96000: sr_code = 11; break; default: sr_code = 0;
    }
    uint32_t bd_code = 0;
    switch (p.bps) {
        case 8: bd_code = 1; break; case 12: bd_code = 2; break; case 16: bd_code = 4; break;
        case 20: bd_code = 5; break; case 24: bd_code = 6; break; case 32: bd_code = 7; break; default: bd_code = 0;
    }
    out.push_back((uint8_t)((bs_code << 4) | sr_code));
    out.push_back((uint8_t)((ch_code << 4) | (bd_code << 1)));
    put_utf8(out, p.variable_blocksize ? f.first_sample : f.index);
    if (bs_extra == 1) out.push_back((uint8_t)(bs - 1));
    else if (bs_extra == 2) { out.push_back((uint8_t)((bs - 1) >> 8)); out.push_back((uint8_t)(bs - 1)); }
    out.push_back(crc8(out.data() + start, out.size() - start));
    // subframes
    BitWriter bw(out);
    const uint32_t acc_bits = cb == 1 ? 16 : cb == 2 ? 32 : 64;
    for (uint32_t c = 0; c < C; ++c) {
        const std::vector<int64_t> *sig = &ch[c];
        if (mode == 1 && c == 1) sig = &b;
        if (mode == 2 && c == 0) sig = &a;
        if (mode == 3) sig = c == 0 ? &a : &b;
        const bool side = c == side_channel;
        encode_subframe(p, bw, *sig, bs, p.bps + (side ? 1 : 0), !side, acc_bits, f.index * 8 + c + (p.seed & 7));
    }
    bw.align();
    const uint16_t crc = crc16(out.data() + start, out.size() - start);
    out.push_back((uint8_t)(crc >> 8));
    out.push_back((uint8_t)crc);
}

int generate(const fg_params &p, int want_pcm, fg_result *out, uint32_t threads) {
    memset(out, 0, sizeof(*out));
    if (p.channels < 1 || p.channels > 8 || p.bps < 4 || p.bps > 32 || p.n_samples == 0) return 1;
    if (!p.variable_blocksize && (p.block_size < 1 || p.block_size > 65535)) return 1;
    if (p.lpc_order > 32 || p.qlp_precision < 1 || p.qlp_precision > 15) return 1;
    const std::vector<FramePlan> plan = plan_frames(p);
    const size_t nf = plan.size();
    const uint32_t cb = container_of(p.bps);
    const uint64_t pcm_bytes = p.n_samples * p.channels * cb;
    uint8_t *pcm = (uint8_t *)aligned_alloc(64, (pcm_bytes + 63) & ~63ull);
    if (!pcm) return 2;
    if (threads == 0) threads = std::max(1u, std::thread::hardware_concurrency());
    threads = (uint32_t)std::min<size_t>(threads, std::max<size_t>(1, nf / 4));
    // frames are encoded in contiguous chunks, several per thread, stitched in order afterwards
    const size_t n_chunks = std::min<size_t>(nf, (size_t)threads * 8);
    std::vector<std::vector<uint8_t>> chunk_bytes(n_chunks);
    std::vector<uint32_t> frame_len(nf);
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        std::vector<std::vector<int64_t>> ch;
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= n_chunks) break;
            const size_t f0 = nf * k / n_chunks, f1 = nf * (k + 1) / n_chunks;
            std::vector<uint8_t> &o = chunk_bytes[k];
            o.reserve((f1 - f0) * (size_t)plan[f0].block_size * p.channels * ((p.bps + 7) / 8) / 2 + 1024);
            for (size_t f = f0; f < f1; ++f) {
                const size_t before = o.size();
                encode_frame(p, plan[f], o, pcm, ch);
                frame_len[f] = (uint32_t)(o.size() - before);
            }
        }
    };
    const auto t_a = std::chrono::steady_clock::now();
    if (threads <= 1) worker();
    else {
        std::vector<std::thread> ts;
        for (uint32_t t = 0; t < threads; ++t) ts.emplace_back(worker);
        for (auto &t : ts) t.join();
    }
    const auto t_b = std::chrono::steady_clock::now();
    // MD5 over the unshifted little-endian samples, bytes per sample = ceil(bps/8) (zflac.zig:267-277, Q4)
    Md5 md5;
    {
        const uint32_t hb = (p.bps + 7) / 8;
        const uint32_t msb = (p.bps >= 9 && p.bps <= 15) ? 16 - p.bps : (p.bps >= 17 && p.bps <= 31) ? 32 - p.bps : 0;
        const uint64_t n = p.n_samples * p.channels;
        if (hb == cb && msb == 0) md5.update(pcm, pcm_bytes);
        else {
            std::vector<uint8_t> tmp;
            tmp.reserve(1 << 20);
            for (uint64_t i = 0; i < n; ++i) {
                int32_t v = cb == 1 ? ((int8_t *)pcm)[i] : cb == 2 ? ((int16_t *)pcm)[i] : ((int32_t *)pcm)[i];
                v >>= msb;
                for (uint32_t k = 0; k < hb; ++k) tmp.push_back((uint8_t)((uint32_t)v >> (8 * k)));
                if (tmp.size() >= (1 << 20) - 8) { md5.update(tmp.data(), tmp.size()); tmp.clear(); }
            }
            md5.update(tmp.data(), tmp.size());
        }
        md5.finish(out->md5);
    }
    const auto t_c = std::chrono::steady_clock::now();
    uint64_t audio = 0;
    uint32_t min_f = 0xFFFFFF, max_f = 0, min_b = 65535, max_b = 0;
    for (size_t f = 0; f < nf; ++f) {
        audio += frame_len[f];
        min_f = std::min(min_f, frame_len[f]);
        max_f = std::max(max_f, frame_len[f]);
        max_b = std::max(max_b, plan[f].block_size);
        if (f + 1 < nf || nf == 1) min_b = std::min(min_b, plan[f].block_size);
    }
    if (max_f > 0xFFFFFF) max_f = 0;
    const uint64_t total_bytes = 4 + 4 + 34 + audio;
    uint8_t *fl = (uint8_t *)aligned_alloc(64, (total_bytes + 64 + 63) & ~63ull);
    if (!fl) { free(pcm); return 2; }
    uint8_t *w = fl;
    memcpy(w, "fLaC", 4); w += 4;
    *w++ = 0x80; *w++ = 0; *w++ = 0; *w++ = 34;  // last block, STREAMINFO, 34 bytes
    *w++ = (uint8_t)(min_b >> 8); *w++ = (uint8_t)min_b;
    *w++ = (uint8_t)(max_b >> 8); *w++ = (uint8_t)max_b;
    *w++ = (uint8_t)(min_f >> 16); *w++ = (uint8_t)(min_f >> 8); *w++ = (uint8_t)min_f;
    *w++ = (uint8_t)(max_f >> 16); *w++ = (uint8_t)(max_f >> 8); *w++ = (uint8_t)max_f;
    const uint64_t total = p.set_total ? p.n_samples : 0;
    const uint64_t packed = ((uint64_t)p.sample_rate << 44) | ((uint64_t)(p.channels - 1) << 41) | ((uint64_t)(p.bps - 1) << 36) | (total & 0xFFFFFFFFFull);
    for (int i = 7; i >= 0; --i) *w++ = (uint8_t)(packed >> (8 * i));
    memcpy(w, out->md5, 16); w += 16;
    for (auto &cbv : chunk_bytes) {
        memcpy(w, cbv.data(), cbv.size());
        w += cbv.size();
        std::vector<uint8_t>().swap(cbv);
    }
    memset(w, 0, 64);
    if (getenv("FLACGEN_TIMING")) {
        const auto t_d = std::chrono::steady_clock::now();
        fprintf(stderr, "flacgen: encode %.3f s (%u threads), md5 %.3f s, stitch %.3f s\n", std::chrono::duration<double>(t_b - t_a).count(), threads,
                std::chrono::duration<double>(t_c - t_b).count(), std::chrono::duration<double>(t_d - t_c).count());
    }
    out->flac = fl;
    out->flac_bytes = total_bytes;
    out->n_frames = nf;
    out->audio_offset = 42;
    out->container_bytes = cb;
    out->pcm_bytes = pcm_bytes;
    if (want_pcm) out->pcm = pcm;
    else { free(pcm); out->pcm = nullptr; }
    return 0;
}

}  // namespace

extern "C" {

int fg_generate(const fg_params *p, int want_pcm, fg_result *out) {
    if (!p || !out || p->struct_size != sizeof(fg_params)) return 1;
    return generate(*p, want_pcm, out, p->n_threads);
}

int fg_generate_batch(const fg_params *ps, uint32_t n, int want_pcm, fg_result *outs, uint32_t n_threads) {
    if (!ps || !outs) return 1;
    if (n_threads == 0) n_threads = std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> next{0};
    std::atomic<int> rc{0};
    auto worker = [&]() {
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= n) break;
            const int e = (ps[i].struct_size == sizeof(fg_params)) ? generate(ps[i], want_pcm, &outs[i], 1) : 1;
            if (e) rc = e;
        }
    };
    std::vector<std::thread> ts;
    for (uint32_t t = 0; t < std::min(n_threads, std::max(1u, n)); ++t) ts.emplace_back(worker);
    for (auto &t : ts) t.join();
    return rc;
}

void fg_free(fg_result *r) {
    if (!r) return;
    free(r->flac);
    free(r->pcm);
    r->flac = nullptr;
    r->pcm = nullptr;
}

}  // extern "C"
