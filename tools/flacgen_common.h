// flacgen_common.h — what the CPU generator (flacgen.cpp) and the CUDA generator (flacgen_cuda.cu) share.
// TEST / BENCH INFRASTRUCTURE, not product code.
//
// The two generators must write the same bytes for the same fg_params. Integer work is order independent; the
// floating-point work (signal synthesis, LPC analysis) is not, so every floating-point expression both sides evaluate
// lives HERE, once, and both compilers are told not to contract a*b+c into an FMA (g++ -ffp-contract=off, nvcc
// -fmad=false): +, -, *, / and the int<->double conversions are then correctly rounded IEEE operations on either
// machine. The transcendental calls (sin, cos, fmod) are not reproducible between glibc and the CUDA math library, so
// they only ever run on the HOST: per frame phases and Hann windows are tables the CUDA generator uploads.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define FG_HD __host__ __device__ inline
#else
#define FG_HD inline
#endif

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
}

namespace fgc {

// ------------------------------------------------------------------ RNG
struct Xo {
    uint64_t s[4];
    FG_HD static uint64_t splitmix(uint64_t &x) {
        uint64_t z = (x += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    FG_HD explicit Xo(uint64_t seed) {
        for (int i = 0; i < 4; ++i) s[i] = splitmix(seed);
    }
    FG_HD static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    FG_HD uint64_t next() {
        const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return r;
    }
    // ~N(0,1): sum of four 16-bit uniforms, centred and scaled (Irwin-Hall, var = 4/12)
    FG_HD double gauss() {
        const uint64_t r = next();
        const double u = (double)(r & 0xFFFF) + (double)((r >> 16) & 0xFFFF) + (double)((r >> 32) & 0xFFFF) + (double)(r >> 48);
        return (u - 2.0 * 65535.0) * (1.0 / (65536.0 * 0.5773502691896258));
    }
};
FG_HD uint64_t frame_rng_seed(uint64_t seed, uint64_t frame_index) {
    return seed * 0x9E3779B97F4A7C15ull + frame_index * 0xD1B54A32D192ED03ull + 1;
}

// ------------------------------------------------------------------ small integer helpers
FG_HD uint32_t container_of(uint32_t bps) { return bps <= 8 ? 1 : bps <= 16 ? 2 : 4; }
FG_HD uint32_t msb_shift_of(uint32_t bps) { return (bps >= 9 && bps <= 15) ? 16 - bps : (bps >= 17 && bps <= 31) ? 32 - bps : 0; }
FG_HD uint64_t zigzag(int64_t r) { return ((uint64_t)r << 1) ^ (uint64_t)(r >> 63); }
FG_HD int bits_needed_signed(int64_t lo, int64_t hi) {  // two's complement width holding [lo, hi]
    int b = 1;
    while (b < 64 && (lo < -(1ll << (b - 1)) || hi > (1ll << (b - 1)) - 1)) ++b;
    return b;
}
FG_HD uint32_t var_block_size(uint64_t idx) {
    const uint32_t k[7] = {16, 192, 576, 1000, 4096, 4608, 16384};
    return k[idx % 7];
}
FG_HD int fixed_coef(uint32_t order, uint32_t j) {
    const int k[5][4] = {{0, 0, 0, 0}, {1, 0, 0, 0}, {2, -1, 0, 0}, {3, -3, 1, 0}, {4, -6, 4, -1}};
    return k[order][j];
}
// channel assignment of a frame: 0 independent, 1 left/side, 2 side/right, 3 mid/side
FG_HD uint32_t stereo_mode_of(const fg_params &p, uint64_t frame_index) {
    if (p.channels != 2) return 0;
    const uint32_t mode = p.stereo_mode == 4 ? (uint32_t)(frame_index % 4) : p.stereo_mode;
    return p.bps == 32 ? 0 : mode;  // a 33-bit side channel does not fit the reference's container (Q9)
}
FG_HD uint32_t wasted_of_frame(const fg_params &p, uint64_t frame_index) {  // low bits cleared on every channel of the frame
    if (!p.wasted_bits) return 0;
    const int w = 1 + (int)(frame_index % 9);
    return w < (int)p.bps - 2 ? (uint32_t)w : 0;
}
// the channel a subframe_mix stream makes constant in this frame (never one that becomes a side channel, Q8), or 99
FG_HD uint32_t constant_channel_of(const fg_params &p, uint64_t frame_index) {
    if (!p.subframe_mix || frame_index % 5 != 2) return 99;
    return (uint32_t)((frame_index / 5) % p.channels);
}

// ------------------------------------------------------------------ frame header
struct FrameHeaderBytes {
    uint8_t b[16];
    uint32_t n;  // including the CRC-8
};
FG_HD uint8_t crc8_update(uint8_t c, uint8_t byte) {
    c ^= byte;
    for (int k = 0; k < 8; ++k) c = (c & 0x80) ? (uint8_t)((c << 1) ^ 0x07) : (uint8_t)(c << 1);
    return c;
}
FG_HD FrameHeaderBytes frame_header(const fg_params &p, uint64_t frame_index, uint64_t first_sample, uint32_t bs) {
    FrameHeaderBytes h;
    uint32_t n = 0;
    const uint32_t mode = stereo_mode_of(p, frame_index);
    const uint32_t ch_code = (p.channels == 2 && mode) ? 7 + mode : p.channels - 1;
    h.b[n++] = 0xFF;
    h.b[n++] = p.variable_blocksize ? 0xF9 : 0xF8;
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
    h.b[n++] = (uint8_t)((bs_code << 4) | sr_code);
    h.b[n++] = (uint8_t)((ch_code << 4) | (bd_code << 1));
    const uint64_t v = p.variable_blocksize ? first_sample : frame_index;
    if (v < 0x80) h.b[n++] = (uint8_t)v;
    else {  // the "UTF-8" number: nb bytes carry 5 nb + 1 bits
        int nb = 2;
        while (nb < 7 && v >= (1ull << (5 * nb + 1))) ++nb;
        h.b[n++] = (uint8_t)((0xFF << (8 - nb)) | (v >> (6 * (nb - 1))));
        for (int i = nb - 2; i >= 0; --i) h.b[n++] = (uint8_t)(0x80 | ((v >> (6 * i)) & 0x3F));
    }
    if (bs_extra == 1) h.b[n++] = (uint8_t)(bs - 1);
    else if (bs_extra == 2) { h.b[n++] = (uint8_t)((bs - 1) >> 8); h.b[n++] = (uint8_t)(bs - 1); }
    uint8_t c = 0;
    for (uint32_t i = 0; i < n; ++i) c = crc8_update(c, h.b[i]);
    h.b[n++] = c;
    h.n = n;
    return h;
}

// ------------------------------------------------------------------ signal
// Three tones by phasor rotation from the absolute sample index (continuous across frames) plus AR(2)-shaped noise;
// the second channel of a stereo stream follows the first (R = 0.8 L + its own noise: gives mid/side something to win).
struct ToneStep {   // per (stream, channel): the rotation of one sample
    double cr[3], ci[3];
};
struct ToneState {  // per (frame, channel): the phasors at the frame's first sample
    double re[3], im[3];
};
struct SignalConsts {  // per stream
    double amp175;  // tone amplitude / 1.75 (the three tones' relative levels sum to 1.75)
    double limit;   // clamp: side must fit the container (Q9)
    double full;
};

// HOST ONLY (no FG_HD): everything that calls libm's sin / cos / fmod / ldexp.
inline double tone_omega(const fg_params &p, uint32_t c, int t) {
    static const double kFreq[3] = {220.0, 1318.5, 5274.0};
    return 2.0 * M_PI * kFreq[t] * (1.0 + 0.013 * c) / (double)p.sample_rate;
}
inline ToneStep tone_step(const fg_params &p, uint32_t c) {
    ToneStep s;
    for (int t = 0; t < 3; ++t) {
        const double w = tone_omega(p, c, t);
        s.cr[t] = std::cos(w);
        s.ci[t] = std::sin(w);
    }
    return s;
}
inline ToneState tone_state(const fg_params &p, uint32_t c, uint64_t first_sample) {
    ToneState s;
    for (int t = 0; t < 3; ++t) {
        const double w = tone_omega(p, c, t);
        const double ph = std::fmod(w * (double)first_sample, 2.0 * M_PI) + 0.7 * c;
        s.re[t] = std::cos(ph);
        s.im[t] = std::sin(ph);
    }
    return s;
}
inline SignalConsts signal_consts(const fg_params &p) {
    SignalConsts k;
    k.full = std::ldexp(1.0, (int)p.bps - 1);
    k.amp175 = p.amplitude * k.full / 1.75;
    k.limit = (p.channels == 2 && p.stereo_mode != 0) ? 0.49 * k.full : 0.98 * k.full;
    return k;
}
inline double frame_sigma(const fg_params &p, const SignalConsts &k, uint64_t frame_index) {
    double sigma = p.noise * k.full;
    if (p.signal == 1) {  // wide dynamic range: noise level sweeps over the frames, tones stay
        const int span = std::max(1, (int)p.bps - 4);
        sigma = std::ldexp(0.6, (int)(frame_index % (uint64_t)span));
        if (sigma > 0.2 * k.full) sigma = 0.2 * k.full;
    }
    return sigma;
}
inline double hann(uint32_t i, uint32_t n) {
    const double t = (double)i / (double)(n - 1 ? n - 1 : 1);
    return 0.5 - 0.5 * std::cos(2.0 * M_PI * t);
}

struct ChannelSynth {  // the per-sample recurrence of one channel of one frame
    ToneState ph;
    double z1, z2;  // AR(2) noise shaping: y = e + 1.2 z1 - 0.5 z2  (spectral tilt)
    double sg_gain;
    FG_HD void init(const ToneState &s, double sigma, bool follow) {
        ph = s;
        z1 = 0;
        z2 = 0;
        const double ar_gain = 0.45;
        const double sg = follow ? sigma * 0.66 : sigma;
        sg_gain = sg * ar_gain;
    }
    // the sample before clamping; `base` is the first channel's (unclamped) sample, used when `follow`
    FG_HD double next(Xo &rng, const ToneStep &st, const SignalConsts &k, bool follow, double base) {
        double v = 0;
        if (!follow) {
            const double rel[3] = {1.0, 0.5, 0.25};
            for (int t = 0; t < 3; ++t) {
                v += rel[t] * ph.im[t];
                const double nr = ph.re[t] * st.cr[t] - ph.im[t] * st.ci[t];
                ph.im[t] = ph.re[t] * st.ci[t] + ph.im[t] * st.cr[t];
                ph.re[t] = nr;
            }
            v *= k.amp175;
        } else {
            v = 0.8 * base;
        }
        const double e = rng.gauss();
        const double y = e + 1.2 * z1 - 0.5 * z2;
        z2 = z1;
        z1 = y;
        v += sg_gain * y;
        return v;
    }
};
FG_HD int64_t quantize_sample(double v, double limit) {
    if (v > limit - 1) v = limit - 1;
    if (v < -limit) v = -limit;
    return (int64_t)llrint(v);
}

// ------------------------------------------------------------------ LPC analysis
// Quantised LPC coefficients from the autocorrelation ac[0..order] of the Hann-windowed block (libFLAC-style:
// Levinson-Durbin, shift chosen so the largest coefficient uses the full precision, error feedback on rounding).
// The autocorrelation itself is a[lag] = sum over i = lag..n-1, IN THAT ORDER, of w[i] * w[i - lag], w[i] = hann(i, n) * s[i].
FG_HD bool lpc_from_autocorr(double *ac, uint32_t order, uint32_t precision, int32_t *qcoef, uint32_t &shift) {
    if (ac[0] <= 0) return false;
    ac[0] *= 1.0 + 1e-9;
    double lpc[33], tmp[33];
    for (int i = 0; i < 33; ++i) lpc[i] = 0;
    double err = ac[0];
    for (uint32_t i = 0; i < order; ++i) {
        double r = -ac[i + 1];
        for (uint32_t j = 0; j < i; ++j) r -= lpc[j] * ac[i - j];
        r /= err;
        for (uint32_t j = 0; j < i; ++j) tmp[j] = lpc[j];
        lpc[i] = r;
        for (uint32_t j = 0; j < i; ++j) lpc[j] = tmp[j] + r * tmp[i - 1 - j];
        err *= (1.0 - r * r);
        if (err <= 0) { err = 1e-9; }
    }
    double cmax = 0;
    for (uint32_t i = 0; i < order; ++i) {
        const double a = fabs(lpc[i]);
        if (a > cmax) cmax = a;
    }
    if (cmax <= 0) return false;
    int log2c;
    frexp(cmax, &log2c);  // cmax = m * 2^log2c, m in [0.5, 1)
    int sh = (int)precision - log2c - 1;
    if (sh > 15) sh = 15;
    if (sh < 0) return false;  // would need a negative shift: not representable for the reference (Q17)
    const int32_t qmax = (1 << (precision - 1)) - 1, qmin = -(1 << (precision - 1));
    double errf = 0;
    for (uint32_t i = 0; i < order; ++i) {
        errf += -lpc[i] * (double)(1 << sh);
        long long q = llrint(errf);
        if (q > qmax) q = qmax;
        if (q < qmin) q = qmin;
        qcoef[i] = (int32_t)q;
        errf -= (double)q;
    }
    shift = (uint32_t)sh;
    return true;
}

// ------------------------------------------------------------------ predictor choice (everything but the LPC fit)
enum : uint32_t { kVerbatim = 0, kFixed = 1, kLpc = 2, kConstant = 3 };
struct Choice {
    uint32_t kind, order, prec;
};
// What encode_subframe tries first for a non-constant subframe of `eff` bits (after the wasted bits went). kLpc may
// still fall back to FIXED min(2, order) when the fit fails (lpc_fallback); kFixed may still become VERBATIM when the
// accumulator is too narrow (fixed_allowed) or a residual leaves +-2^30.
FG_HD Choice choose_predictor(const fg_params &p, uint64_t salt, uint32_t bs, uint32_t eff, uint32_t acc_bits) {
    Choice c;
    c.kind = kLpc;
    c.order = p.lpc_order;
    c.prec = p.qlp_precision;
    if (p.subframe_mix) {
        const uint64_t sel = salt % 8;
        if (sel == 7) c.kind = kVerbatim;
        else if (sel >= 2 && sel <= 6) { c.kind = kFixed; c.order = (uint32_t)(sel - 2); }
        else if (sel == 1) c.order = 1 + (uint32_t)((salt / 8) % p.lpc_order);
    }
    if (c.order > 32) c.order = 32;
    if (c.kind == kLpc && (c.order == 0 || bs <= c.order)) c.kind = kVerbatim;
    if (c.kind == kFixed && bs <= c.order) c.kind = kVerbatim;
    if (c.kind == kLpc) {
        // keep |sum| <= 2^(eff-1 + prec-1 + ceil(log2(order))) inside the reference's accumulator (Q10)
        uint32_t lg = 0;
        while ((1u << lg) < c.order) ++lg;
        while (c.prec > 1 && eff + c.prec + lg > acc_bits) --c.prec;
        if (eff + c.prec + lg > acc_bits) { c.kind = kFixed; c.order = c.order < 2 ? c.order : 2; }
    }
    return c;
}
FG_HD bool fixed_allowed(uint32_t eff, uint32_t acc_bits) { return !(eff + 4 > acc_bits - 1); }
FG_HD uint64_t subframe_salt(const fg_params &p, uint64_t frame_index, uint32_t c) { return frame_index * 8 + c + (p.seed & 7); }

// ------------------------------------------------------------------ Rice partitions
FG_HD uint32_t rice_kmax(const fg_params &p) { return p.rice_method ? 30 : 14; }
FG_HD uint32_t rice_pbits(const fg_params &p) { return p.rice_method ? 5 : 4; }
FG_HD bool partition_order_valid(uint32_t bs, uint32_t po, uint32_t pred_order) {
    // must divide the block; first partition may not be negative (Q16)
    return ((bs >> po) << po) == bs && (bs >> po) >= pred_order && (bs >> po) > 0;
}
// the range [lo, hi] of partition orders the planner walks (it stops at the first invalid one)
FG_HD void partition_order_range(const fg_params &p, uint32_t bs, uint32_t pred_order, uint32_t &lo, uint32_t &hi) {
    lo = p.min_partition_order;
    hi = p.max_partition_order < 15 ? p.max_partition_order : 15;
    while (lo > 0 && !partition_order_valid(bs, lo, pred_order)) --lo;  // short last block: the largest valid order below
}
FG_HD uint32_t rice_k_of_mean(uint64_t sum, uint32_t cnt) {  // k ~ floor(log2(mean)); k + 1 is tried too
    uint32_t k = 0;
    if (cnt) {
        const uint64_t mean = sum / cnt;
        while (k < 40 && (mean >> (k + 1)) != 0) ++k;
    }
    return k;
}
FG_HD bool partition_forced_escape(const fg_params &p, uint64_t salt, uint32_t q) {
    return p.escape_every && ((salt + q) % p.escape_every == p.escape_every - 1);
}

// ------------------------------------------------------------------ CRC-16 (poly 0x8005, MSB first), slice-by-8 tables
FG_HD void crc16_make_tables(uint16_t (*t)[256]) {
    for (int i = 0; i < 256; ++i) {
        uint32_t d = (uint32_t)i << 8;
        for (int k = 0; k < 8; ++k) d = (d & 0x8000) ? ((d << 1) ^ 0x8005) & 0xFFFF : (d << 1) & 0xFFFF;
        t[0][i] = (uint16_t)d;
    }
    for (int s = 1; s < 8; ++s)
        for (int i = 0; i < 256; ++i) {
            const uint16_t v = t[s - 1][i];
            t[s][i] = (uint16_t)((v << 8) ^ t[0][v >> 8]);
        }
}

// ------------------------------------------------------------------ MD5 (RFC 1321)
struct Md5 {
    uint32_t a = 0x67452301u, b = 0xefcdab89u, c = 0x98badcfeu, d = 0x10325476u;
    uint64_t total = 0;
    uint8_t buf[64];
    size_t fill = 0;
    FG_HD static uint32_t rl(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }
    FG_HD void block(const uint8_t *p) {
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
    FG_HD void update(const uint8_t *p, size_t n) {
        total += n;
        if (fill) {
            const size_t k = n < 64 - fill ? n : 64 - fill;
            memcpy(buf + fill, p, k);
            fill += k; p += k; n -= k;
            if (fill == 64) { block(buf); fill = 0; }
        }
        for (; n >= 64; p += 64, n -= 64) block(p);
        if (n) { memcpy(buf, p, n); fill = n; }
    }
    FG_HD void finish(uint8_t out[16]) {
        const uint64_t bits = total * 8;
        uint8_t pad[72];
        for (int i = 0; i < 72; ++i) pad[i] = 0;
        pad[0] = 0x80;
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

}  // namespace fgc
