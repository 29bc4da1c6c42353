// predict.cu — K3p predict_epilogue (sm_100a): pass B of the two-pass subframe decoder fused with
// the epilogue. One lane per SUBFRAME runs the predictor recurrence the reference runs inside its
// frame loop — FIXED (src/zflac.zig:466-498), LPC (:499-541, :604-612), warm-up samples kept as
// they are, CONSTANT broadcast (:445-454), VERBATIM copied (:455-465) — over the residuals pass A
// (parse.cu) left in the planar scratch, then applies the wasted-bits shift (:494, :537), the
// stereo decorrelation (:553-578; the partner channel is the neighbouring lane: one shuffle), the
// narrowing store into the container type and the MSB-align shift (:287-306), and writes the
// interleaved frame-major / channel-minor PCM (:443).
//
// Mapping. CP = channels rounded up to a power of two lanes per frame, G = 32 / CP frames per warp.
// Coefficients and the last P samples live in registers (P = 4 / 8 / 12 / 32 taps, chosen per warp
// from the largest order among its lanes; int32 sums unless a lane's SubRec says the sum needs 64
// bits). Eight time steps per iteration: 2 x 128-bit residual loads per lane (prefetched one
// iteration ahead), 8 x P IMADs, then the 8 x C container values of every frame are staged as
// their PCM byte image in shared memory and leave as 16-byte stores.
#include <type_traits>

#include "kernels.cuh"
#include "parse_core.cuh"

namespace zf {

constexpr int kPredThreads = 128;
constexpr int kPredWarps = kPredThreads / 32;
constexpr int kTileBytes = 2 * (1024 + 32 * 16);  // double buffered; per buffer: sum over a warp's frames of (8*C*CB + 16) <= 1024 + 512
static_assert(sizeof(SubRec) == kSubRecBytes, "SubRec size");

struct PredLane {           // what a lane knows about its subframe
    const int32_t *src;     // its scratch plane
    uint32_t bs;            // block size (0: idle lane)
    uint32_t order, shift, wasted, kind;
    int32_t constant;
    uint32_t mode;          // decorrelation role: 0 none, 1 out = p - s, 2 out = s + p, 3 mid/side ch0, 4 mid/side ch1
    uint64_t elem_base;     // element index of (row 0, this channel) in the PCM buffer
    uint32_t C;             // channels of the frame
    uint32_t staged_ok;     // the frame's PCM address allows the staged vector path
};

__device__ __forceinline__ int32_t narrow(int32_t v, uint32_t cb) {  // @intCast narrowing (wraps where the reference is undefined)
    return cb == 1 ? (int32_t)(int8_t)v : cb == 2 ? (int32_t)(int16_t)v : v;
}

template <int P, typename ACC>
__device__ __forceinline__ void predict_run(const PredLane &L, const int32_t (&cf)[P], uint32_t bs_max, uint32_t cb,
                                            uint32_t msb_shift, uint32_t cp_shift, bool stereo, bool wide_pair,
                                            uint8_t *__restrict__ pcm, uint8_t *tile, uint32_t unit, uint32_t n_pieces,
                                            const uint64_t (&rd_base)[2], const uint32_t (&rd_bs)[2], const uint32_t (&rd_soff)[2],
                                            const uint32_t (&rd_step)[2], uint32_t &lossy) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t g = lane >> cp_shift, c = lane & ((1u << cp_shift) - 1u);
    const uint32_t chunk_bytes = 8 * L.C * cb;
    int32_t h[P];
#pragma unroll
    for (int j = 0; j < P; ++j) h[j] = 0;
    const bool loads = L.bs != 0 && L.kind != kSubConst;
    int4 na = make_int4(L.constant, L.constant, L.constant, L.constant), nb = na;
    if (loads) {
        na = __ldcs(reinterpret_cast<const int4 *>(L.src));
        nb = __ldcs(reinterpret_cast<const int4 *>(L.src + 4));
    }
    uint32_t buf = 0;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < bs_max; i0 += 8) {
        const int4 a = na, b = nb;
        if (loads && i0 + 8 < L.bs) {  // next chunk (the plane pitch is a multiple of 8: whole chunks are in bounds)
            na = __ldcs(reinterpret_cast<const int4 *>(L.src + i0 + 8));
            nb = __ldcs(reinterpret_cast<const int4 *>(L.src + i0 + 12));
        }
        const int32_t r[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        int32_t x[P + 8];
#pragma unroll
        for (int j = 0; j < P; ++j) x[j] = h[j];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            ACC acc = 0;
#pragma unroll
            for (int j = P - 1; j >= 0; --j) acc += (ACC)cf[j] * (ACC)x[P + u - 1 - j];  // newest tap last
            const int32_t pred = (int32_t)(acc >> L.shift);
            x[P + u] = r[u] + (i0 + u >= L.order ? pred : 0);  // warm-up samples stay as read
        }
#pragma unroll
        for (int j = 0; j < P; ++j) h[j] = x[j + 8];

        // ---- wasted bits, decorrelation, narrowing, MSB-align shift
        int32_t o[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int32_t s = (int32_t)((uint32_t)x[P + u] << L.wasted);
            int32_t v = s;
            if (stereo) {  // warp-uniform: every lane takes part in the shuffle
                const int32_t p = __shfl_xor_sync(0xFFFFFFFFu, s, 1);
                if (!wide_pair) {
                    if (L.mode == 1) v = p - s;
                    else if (L.mode == 2) v = s + p;
                    else if (L.mode == 3) v = (int32_t)(((uint32_t)s << 1) + (uint32_t)(p & 1) + (uint32_t)p) >> 1;
                    else if (L.mode == 4) v = (int32_t)(((uint32_t)p << 1) + (uint32_t)(s & 1) - (uint32_t)s) >> 1;
                } else {  // 30/31-bit streams: the mid/side sum needs more than 32 bits
                    const int64_t S = s, Q = p;
                    if (L.mode == 1) v = (int32_t)(Q - S);
                    else if (L.mode == 2) v = (int32_t)(S + Q);
                    else if (L.mode == 3) v = (int32_t)((((S << 1) + (Q & 1)) + Q) >> 1);
                    else if (L.mode == 4) v = (int32_t)((((Q << 1) + (S & 1)) - S) >> 1);
                }
            }
            const int32_t cv = narrow(v, cb);
            const int32_t sh = narrow((int32_t)((uint32_t)cv << msb_shift), cb);  // `<<=` discards the bits shifted out
            if ((sh >> msb_shift) != cv && i0 + u < L.bs) lossy = 1;
            o[u] = sh;
        }

        // ---- store: staged PCM byte image -> 16-byte stores, or element by element (tails, odd addresses)
        const bool staged = L.staged_ok && i0 + 8 <= L.bs;
        uint8_t *tb = tile + buf * (kTileBytes / 2);
        if (L.bs != 0) {
            if (staged) {
                uint8_t *slot = tb + g * (chunk_bytes + 16) + c * cb;
                const uint32_t rs = L.C * cb;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (cb == 2) *reinterpret_cast<int16_t *>(slot + u * rs) = (int16_t)o[u];
                    else if (cb == 4) *reinterpret_cast<int32_t *>(slot + u * rs) = o[u];
                    else *reinterpret_cast<int8_t *>(slot + u * rs) = (int8_t)o[u];
                }
            } else if (i0 < L.bs) {
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (i0 + u < L.bs) {
                        const uint64_t e = L.elem_base + (uint64_t)(i0 + u) * L.C;
                        if (cb == 2) reinterpret_cast<int16_t *>(pcm)[e] = (int16_t)o[u];
                        else if (cb == 4) reinterpret_cast<int32_t *>(pcm)[e] = o[u];
                        else reinterpret_cast<int8_t *>(pcm)[e] = (int8_t)o[u];
                    }
                }
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (lane + 32u * k < n_pieces && i0 + 8 <= rd_bs[k]) {  // rd_bs = 0 where the frame is not staged
                const uint8_t *sp = tb + rd_soff[k];
                uint8_t *gp = pcm + rd_base[k] + (uint64_t)(i0 >> 3) * rd_step[k];
                if (unit == 16) *reinterpret_cast<uint4 *>(gp) = *reinterpret_cast<const uint4 *>(sp);
                else *reinterpret_cast<uint2 *>(gp) = *reinterpret_cast<const uint2 *>(sp);
            }
        }
        buf ^= 1u;  // the next iteration's stores go to the other half: one __syncwarp per iteration is enough
    }
}

// LARGE = false: warps whose largest order is <= 12 (register windows of 4 / 8 / 12 taps, 4 CTAs per SM);
// LARGE = true: the others (32 taps, twice the registers). Both kernels run over all frames and a warp
// leaves at once when it belongs to the other one; `big_orders` (set by pass A) lets the LARGE launch
// retire immediately when no subframe of the launch has an order above 12.
template <bool LARGE>
__global__ void __launch_bounds__(kPredThreads, LARGE ? 2 : 4)
k_predict_epilogue(const FrameRec *__restrict__ frames, const FrameOut *__restrict__ outs, uint32_t n_frames,
                   const int32_t *__restrict__ scratch, const SubRec *__restrict__ subs, uint8_t *__restrict__ pcm,
                   uint32_t cb, uint32_t msb_shift, uint32_t cp_shift, uint32_t *__restrict__ lossy_flag,
                   const uint32_t *__restrict__ big_orders) {
    __shared__ __align__(16) uint8_t tiles[kPredWarps][kTileBytes];
    if (LARGE && *big_orders == 0) return;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = blockIdx.x * kPredWarps + (threadIdx.x >> 5);
    const uint32_t G = 32u >> cp_shift;
    const uint32_t g = lane >> cp_shift, c = lane & ((1u << cp_shift) - 1u);
    const uint64_t f64 = (uint64_t)warp * G + g;
    if ((uint64_t)warp * G >= n_frames) return;  // whole warp idle

    // ---- the lane's subframe
    PredLane L;
    L.src = scratch;
    L.bs = 0;
    L.order = L.shift = L.wasted = 0;
    L.kind = kSubRaw;
    L.constant = 0;
    L.mode = 0;
    L.elem_base = 0;
    L.C = 1;
    L.staged_ok = 0;
    uint32_t wide_acc = 0, bps = 0;
    uint64_t frame_byte_base = 0;
    const SubRec *sub = nullptr;
    if (f64 < n_frames) {
        const uint32_t f = (uint32_t)f64;
        const FrameRec fr = frames[f];
        if (fr.block_size != 0 && outs[f].status == ST_OK && c < fr.channels) {
            sub = subs + (size_t)f * 8 + c;
            const uint4 hd = *reinterpret_cast<const uint4 *>(sub);  // kind, order, shift, wasted | wide_acc | constant
            L.kind = hd.x & 0xFFu;
            L.order = (hd.x >> 8) & 0xFFu;
            L.shift = (hd.x >> 16) & 0xFFu;
            L.wasted = (hd.x >> 24) & 0xFFu;
            wide_acc = hd.y & 0xFFu;
            L.constant = (int32_t)hd.z;
            if (L.order > 32) L.order = 32;
            L.bs = fr.block_size;
            L.C = fr.channels;
            L.src = scratch + fr.scratch_base + (uint64_t)c * fr.scratch_pitch;
            L.elem_base = fr.sample_base + c;
            bps = fr.bps;
            if (fr.channels == 2) {
                if (fr.ch_code == 8) L.mode = c == 1 ? 1 : 0;
                else if (fr.ch_code == 9) L.mode = c == 0 ? 2 : 0;
                else if (fr.ch_code == 10) L.mode = c == 0 ? 3 : 4;
            }
            frame_byte_base = fr.sample_base * cb;
        }
    }
    // ---- warp-uniform shape
    const uint32_t bs_max = __reduce_max_sync(0xFFFFFFFFu, L.bs);
    if (bs_max == 0) return;
    const uint32_t order_max = __reduce_max_sync(0xFFFFFFFFu, L.bs ? L.order : 0u);
    if ((order_max > 12) != LARGE) return;  // the other kernel's warp
    const bool any_wide = __any_sync(0xFFFFFFFFu, L.bs != 0 && wide_acc != 0);
    const uint32_t c_max = __reduce_max_sync(0xFFFFFFFFu, L.bs ? L.C : 0u);
    const uint32_t c_min = __reduce_min_sync(0xFFFFFFFFu, L.bs ? L.C : 0xFFu);
    const bool stereo = __any_sync(0xFFFFFFFFu, L.mode != 0);  // partner channel = lane ^ 1 whatever the group width
    const bool wide_pair = __any_sync(0xFFFFFFFFu, L.mode != 0 && bps > 28);

    // ---- staged store plan: only when every frame of the warp has the same channel count (always, within a stream)
    const uint32_t chunk_bytes = 8 * c_max * cb;
    const uint32_t unit = (chunk_bytes & 15u) == 0 ? 16u : 8u;
    const bool uniform_c = c_max == c_min;
    if (L.bs != 0 && uniform_c) L.staged_ok = ((reinterpret_cast<uintptr_t>(pcm) + frame_byte_base) & (unit - 1)) == 0 ? 1u : 0u;
    const uint32_t ppf = chunk_bytes / unit;  // pieces per frame chunk
    const uint32_t n_pieces = uniform_c ? G * ppf : 0u;
    uint64_t rd_base[2] = {0, 0};
    uint32_t rd_bs[2] = {0, 0}, rd_soff[2] = {0, 0}, rd_step[2] = {0, 0};
#pragma unroll
    for (int k = 0; k < 2; ++k) {  // this lane's reader role: piece p of the warp's staged image
        const uint32_t p = lane + 32u * k;
        const uint32_t pg = ppf ? p / ppf : 0u, pi = ppf ? p % ppf : 0u;
        const uint32_t srcl = (pg << cp_shift) & 31u;
        const uint64_t fb = __shfl_sync(0xFFFFFFFFu, frame_byte_base, srcl);
        const uint32_t fbs = __shfl_sync(0xFFFFFFFFu, L.staged_ok ? L.bs : 0u, srcl);
        if (p < n_pieces) {
            rd_base[k] = fb + (uint64_t)pi * unit;
            rd_bs[k] = fbs;
            rd_soff[k] = pg * (chunk_bytes + 16) + pi * unit;
            rd_step[k] = chunk_bytes;
        }
    }

    uint32_t lossy = 0;
    uint8_t *tile = tiles[threadIdx.x >> 5];
#define ZF_PRED(PP, ACC)                                                                                                        \
    do {                                                                                                                        \
        int32_t cf[PP];                                                                                                         \
        _Pragma("unroll") for (int j = 0; j < PP; ++j) cf[j] = (sub != nullptr && L.kind == kSubPred && (uint32_t)j < L.order) ? (int32_t)sub->coef[j] : 0; \
        predict_run<PP, ACC>(L, cf, bs_max, cb, msb_shift, cp_shift, stereo, wide_pair, pcm, tile, unit, n_pieces, rd_base, rd_bs, \
                             rd_soff, rd_step, lossy);                                                                          \
    } while (0)
    if (LARGE) {
        if (!any_wide) ZF_PRED(32, int32_t);
        else ZF_PRED(32, int64_t);
    } else if (!any_wide) {
        if (order_max <= 4) ZF_PRED(4, int32_t);
        else if (order_max <= 8) ZF_PRED(8, int32_t);
        else ZF_PRED(12, int32_t);
    } else {
        if (order_max <= 8) ZF_PRED(8, int64_t);
        else ZF_PRED(12, int64_t);
    }
#undef ZF_PRED
    if (lossy) atomicOr(lossy_flag, 1u);
}

cudaError_t launch_predict_epilogue(const PredictParams &p, cudaStream_t st) {
    if (p.n_frames == 0) return cudaSuccess;
    uint32_t cp_shift = 0;
    while ((1u << cp_shift) < p.max_channels) ++cp_shift;  // lanes per frame: channels rounded up to a power of two
    if (cp_shift > 3) cp_shift = 3;
    const uint32_t G = 32u >> cp_shift;
    const uint64_t warps = ((uint64_t)p.n_frames + G - 1) / G;
    const uint32_t blocks = (uint32_t)((warps + kPredWarps - 1) / kPredWarps);
    ZF_COUNT_LAUNCH();
    k_predict_epilogue<false><<<blocks, kPredThreads, 0, st>>>(p.frames, p.outs, p.n_frames, (const int32_t *)p.scratch,
                                                               (const SubRec *)p.subs, (uint8_t *)p.pcm, p.container_bytes, p.msb_shift,
                                                               cp_shift, p.lossy_flag, p.big_orders);
    ZF_COUNT_LAUNCH();
    k_predict_epilogue<true><<<blocks, kPredThreads, 0, st>>>(p.frames, p.outs, p.n_frames, (const int32_t *)p.scratch,
                                                              (const SubRec *)p.subs, (uint8_t *)p.pcm, p.container_bytes, p.msb_shift,
                                                              cp_shift, p.lossy_flag, p.big_orders);
    return cudaGetLastError();
}

}  // namespace zf
