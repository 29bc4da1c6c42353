// predict.cu — K3p predict_epilogue (sm_100a): pass B of the two-pass subframe decoder fused with
// the epilogue. One lane per SUBFRAME runs the predictor recurrence the reference runs inside its
// frame loop — FIXED (src/zflac.zig:466-498), LPC (:499-541, :604-612), warm-up samples kept as
// they are, CONSTANT broadcast (:445-454), VERBATIM copied (:455-465) — over the residuals pass A
// (parse.cu) left in the scratch, then applies the wasted-bits shift (:494, :537), the stereo
// decorrelation (:553-578; the partner channel is the neighbouring lane: one shuffle), the narrowing
// store into the container type and the MSB-align shift (:287-306), and writes the interleaved
// frame-major / channel-minor PCM (:443).
//
// Mapping. CP = channels rounded up to a power of two lanes per frame, G = 32 / CP frames per warp,
// one warp per CTA (no block-level cooperation; small CTAs let one wave hold a whole 1-hour stream).
// Coefficients and the last samples live in registers; the tap count of a step (4 / 8 / 12, or 32 in
// the LARGE kernel) and the accumulator width (int32 unless a lane's SubRec says the sum needs 64
// bits) are chosen per warp. Eight time steps per iteration: 2 x 128-bit residual loads per lane
// (prefetched one iteration ahead; the C lanes of a frame read C adjacent sectors, see sub_idx()),
// 8 x P IMADs, then the 8 x C container values of every frame are staged as their PCM byte image in
// shared memory and leave as 16-byte stores.
#include <type_traits>

#include "kernels.cuh"
#include "parse_core.cuh"
#include "parse_narrow.cuh"

namespace zf {

constexpr int kTileBytes = 2 * (1024 + 32 * 16);  // double buffered; per buffer: sum over a warp's frames of (8*C*CB + 16) <= 1024 + 512
static_assert(sizeof(SubRec) == kSubRecBytes, "SubRec size");

template <int CB> __device__ __forceinline__ int32_t narrow(int32_t v) {  // @intCast narrowing (wraps where the reference is undefined)
    return CB == 1 ? (int32_t)(int8_t)v : CB == 2 ? (int32_t)(int16_t)v : v;
}

// Eight steps of the recurrence. h[HMAX-1] is the newest sample; cf[j] multiplies s[i-1-j]; only the
// P newest history slots / first P coefficients take part. WARM: steps below `order` keep the value read.
template <int HMAX, int P, typename ACC, bool WARM>
__device__ __forceinline__ void lpc_chunk(int32_t (&h)[HMAX], const int32_t (&cf)[HMAX], const int32_t (&r)[8], uint32_t shift,
                                          uint32_t order, uint32_t i0, int32_t (&y)[8]) {
    int32_t x[P + 8];
#pragma unroll
    for (int j = 0; j < P; ++j) x[j] = h[HMAX - P + j];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        ACC acc = 0;
#pragma unroll
        for (int j = P - 1; j >= 0; --j) acc += (ACC)cf[j] * (ACC)x[P + u - 1 - j];  // newest tap last
        const int32_t pred = (int32_t)(acc >> shift);
        x[P + u] = r[u] + (WARM ? (i0 + u >= order ? pred : 0) : pred);
        y[u] = x[P + u];
    }
    if (P > 8 || WARM) {  // a window of <= 8 taps is the last eight samples: the older slots are never read again (the warm-up chunk, which
                          // reads all of them, only runs while they are still zero)
#pragma unroll
        for (int j = 0; j < HMAX - 8; ++j) h[j] = h[j + 8];
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) h[HMAX - 8 + u] = y[u];
}

// LARGE = false: warps whose largest order is <= 12 (history of 12 registers, <= 120 registers per
// thread: 17 warps per SM); LARGE = true: the others (32 taps). Both kernels run over all frames and
// a warp leaves at once when it belongs to the other one; `big_orders` (set by pass A) lets the LARGE
// launch retire immediately when no subframe of the launch has an order above 12.
// NARROW: the scratch holds int16 values (parse_narrow.cuh): one 16-byte load per eight steps.
template <bool LARGE, int CB, bool NARROW>
__global__ void __launch_bounds__(32, LARGE ? 8 : 17)
k_predict_epilogue(const FrameRec *__restrict__ frames, const FrameOut *__restrict__ outs, uint32_t n_frames,
                   const int32_t *__restrict__ scratch, const SubRec *__restrict__ subs, uint8_t *__restrict__ pcm,
                   uint32_t msb_shift, uint32_t cp_shift, uint32_t *__restrict__ lossy_flag,
                   const uint32_t *__restrict__ big_orders) {
    constexpr int HMAX = LARGE ? 32 : 12;
    __shared__ __align__(16) uint8_t tile[kTileBytes];
    if (LARGE && *big_orders == 0) return;
    const uint32_t lane = threadIdx.x;
    const uint32_t warp = blockIdx.x;
    const uint32_t G = 32u >> cp_shift;
    const uint32_t g = lane >> cp_shift, c = lane & ((1u << cp_shift) - 1u);
    const uint64_t f64 = (uint64_t)warp * G + g;

    // ---- the lane's subframe
    const uint8_t *row0 = reinterpret_cast<const uint8_t *>(scratch);  // NARROW: row t at row0 + (t >> row_shift) * row_stride + (t & row_mask) * 16
    uint32_t row_shift = 0, row_mask = 0, row_stride = 0, half_off = 16;
    uint32_t bs = 0, order = 0, shift = 0, wasted = 0, kind = kSubRaw, C = 1, wide_acc = 0, bps = 0, mode = 0;
    int32_t constant = 0;
    uint32_t pair_code = 1;  // channel assignment of a two-channel frame (1: independent)
    uint64_t elem_base = 0, frame_byte_base = 0;
    const SubRec *sub = nullptr;
    if (f64 < n_frames) {
        const uint32_t f = (uint32_t)f64;
        const FrameRec fr = frames[f];
        if (fr.block_size != 0 && outs[f].status == ST_OK && c < fr.channels) {
            sub = subs + (size_t)f * 8 + c;
            const uint4 hd = *reinterpret_cast<const uint4 *>(sub);  // kind, order, shift, wasted | wide_acc | constant
            kind = hd.x & 0xFFu;
            order = (hd.x >> 8) & 0xFFu;
            shift = (hd.x >> 16) & 0xFFu;
            wasted = (hd.x >> 24) & 0xFFu;
            wide_acc = hd.y & 0xFFu;
            // a 32-bit container: the reference sums in i64 (src/zflac.zig:314-319) and stays defined when a malformed stream
            // carries samples wider than the frame's depth, which pass A's rule (depth + precision + log2(order) <= 32)
            // assumes they are not; without a shift the low 32 bits of the sum are exact either way
            if (CB == 4 && kind == kSubPred && shift != 0) wide_acc = 1;
            constant = (int32_t)hd.z;
            if (order > 32) order = 32;
            if (kind != kSubPred) order = 0;
            bs = fr.block_size;
            C = fr.channels;
            // sub_idx() layout: 8 values of this channel per group of 8 * C (elements: int32, or uint16 when NARROW)
            // rows of eight values, laid out per warp of pass A or per frame (parse_narrow.cuh, flag in SubRec.pad[0]): int16
            // rows when NARROW, else int32 rows in two 16-byte halves `half_off` apart (the per-frame layout is sub_idx()'s)
            if (((hd.y >> 8) & 0xFFu) == kLayoutWarp) {
                const uint64_t base0 = frames[f & ~31u].scratch_base;
                row_stride = NARROW ? 512u : 1024u;
                row0 = reinterpret_cast<const uint8_t *>(scratch + base0) + (size_t)c * (hd.y >> 16) * row_stride + (f & 31u) * 16u;
                row_shift = 0u;
                row_mask = 0u;
                half_off = 512u;
            } else {
                row0 = reinterpret_cast<const uint8_t *>(scratch + fr.scratch_base) + c * 32u;
                row_shift = NARROW ? 1u : 0u;
                row_mask = NARROW ? 1u : 0u;
                row_stride = 32u * fr.channels;
                half_off = 16u;
            }
            elem_base = fr.sample_base + c;
            bps = fr.bps;
            if (fr.channels == 2) {  // :553-578
                if (fr.ch_code == 8) mode = c == 1 ? 1 : 0;
                else if (fr.ch_code == 9) mode = c == 0 ? 2 : 0;
                else if (fr.ch_code == 10) mode = c == 0 ? 3 : 4;
            }
            frame_byte_base = fr.sample_base * CB;
            if (fr.channels == 2) pair_code = fr.ch_code;
        }
    }
    // ---- warp-uniform shape
    const uint32_t bs_max = __reduce_max_sync(0xFFFFFFFFu, bs);
    if (bs_max == 0) return;
    const uint32_t order_max = __reduce_max_sync(0xFFFFFFFFu, bs ? order : 0u);
    if ((order_max > 12) != LARGE) return;  // the other kernel's warp
    const bool any_wide = __any_sync(0xFFFFFFFFu, bs != 0 && wide_acc != 0);
    const uint32_t c_max = __reduce_max_sync(0xFFFFFFFFu, bs ? C : 0u);
    const uint32_t c_min = __reduce_min_sync(0xFFFFFFFFu, bs ? C : 0xFFu);
    const bool stereo = __any_sync(0xFFFFFFFFu, mode != 0);  // partner channel = lane ^ 1 whatever the group width
    const bool wide_pair = __any_sync(0xFFFFFFFFu, mode != 0 && bps > 28);
    // decorrelation as one branch-free form: v = (da * s + db * p + lsb) >> dsh, s = own value, p = partner's
    //   mode 1 (left/side, ch1): p - s;  mode 2 (side/right, ch0): s + p;
    //   mode 3 (mid/side, ch0): (2s + (p & 1) + p) >> 1;  mode 4 (mid/side, ch1): (2p + (s & 1) - s) >> 1
    const int32_t da = mode == 1 || mode == 4 ? -1 : mode == 3 ? 2 : 1;
    const int32_t db = mode == 0 ? 0 : mode == 4 ? 2 : 1;
    const int32_t mp = mode == 3 ? 1 : 0, ms = mode == 4 ? 1 : 0;
    const uint32_t dsh = mode >= 3 ? 1u : 0u;

    // ---- staged store plan: only when every frame of the warp has the same channel count (always, within a stream)
    const uint32_t chunk_bytes = 8 * c_max * CB;
    const uint32_t unit = (chunk_bytes & 15u) == 0 ? 16u : 8u;
    const bool uniform_c = c_max == c_min;
    const bool staged_ok = bs != 0 && uniform_c && ((reinterpret_cast<uintptr_t>(pcm) + frame_byte_base) & (unit - 1)) == 0;
    const uint32_t ppf = chunk_bytes / unit;  // pieces per frame chunk
    const uint32_t n_pieces = uniform_c ? G * ppf : 0u;
    uint8_t *rd_ptr[2] = {pcm, pcm};
    uint32_t rd_bs[2] = {0, 0}, rd_soff[2] = {0, 0};
#pragma unroll
    for (int k = 0; k < 2; ++k) {  // this lane's reader role: piece p of the warp's staged image
        const uint32_t p = lane + 32u * k;
        const uint32_t pg = ppf ? p / ppf : 0u, pi = ppf ? p % ppf : 0u;
        const uint32_t srcl = (pg << cp_shift) & 31u;
        const uint64_t fb = __shfl_sync(0xFFFFFFFFu, frame_byte_base, srcl);
        const uint32_t fbs = __shfl_sync(0xFFFFFFFFu, staged_ok ? bs : 0u, srcl);
        if (p < n_pieces) {
            rd_ptr[k] = pcm + fb + (uint64_t)pi * unit;
            rd_bs[k] = fbs;  // 0 where the frame is not staged
            rd_soff[k] = pg * (chunk_bytes + 16) + pi * unit;
        }
    }
    const bool two_pieces = n_pieces > 32;

    // ---- 16-bit two-channel frames (CB == 2, every frame of the warp, PCM 16-byte aligned): no shared memory at all. The
    // two lanes of a frame trade HALF a group — the even lane takes time steps 0..3 of both channels, the odd lane 4..7 —
    // and each one undoes the channel assignment for its four steps as (left, right) pairs, which are then exactly the
    // 16 bytes it stores: four shuffles, four packs and one store instead of eight shuffles, eight 16-bit shared stores,
    // a warp barrier and a shared load (r12: 122 -> ~50 of the loop's ~265 instructions).
    //   left = (al * a + bl * b + (b & pm)) >> dsh, right = (ar * a + br * b + (b & pm)) >> dsh, a / b = channel 0 / 1 (:553-578)
    const bool odd = (lane & 1u) != 0;
    const bool pair16 = CB == 2 && cp_shift == 1 && uniform_c && c_max == 2 &&
                        __all_sync(0xFFFFFFFFu, bs == 0 || ((reinterpret_cast<uintptr_t>(pcm) + frame_byte_base) & 15u) == 0);
    const int32_t al = pair_code == 10 ? 2 : 1, bl = pair_code == 9 || pair_code == 10 ? 1 : 0;
    const int32_t ar = pair_code == 10 ? 2 : pair_code == 8 ? 1 : 0, br = pair_code == 8 || pair_code == 10 ? -1 : 1;
    const int32_t kxl = odd ? bl : al, kyl = odd ? al : bl, kxr = odd ? br : ar, kyr = odd ? ar : br;  // x: own value, y: the partner's
    const int32_t pm = pair_code == 10 ? 1 : 0;
    const uint32_t psh = pair_code == 10 ? 1u : 0u;
    uint8_t *const pair_ptr = pcm + frame_byte_base + (odd ? 16u : 0u);

    // ---- coefficients and history
    int32_t cf[HMAX], h[HMAX];
#pragma unroll
    for (int j = 0; j < HMAX; ++j) {
        cf[j] = (sub != nullptr && (uint32_t)j < order) ? (int32_t)sub->coef[j] : 0;
        h[j] = 0;
    }
    const uint32_t pclass = LARGE ? (any_wide ? 1u : 0u)
                                  : any_wide ? (order_max <= 8 ? 3u : 4u) : (order_max <= 4 ? 0u : order_max <= 8 ? 1u : 2u);

    const bool any_wasted = __any_sync(0xFFFFFFFFu, wasted != 0);
    const bool loads = bs != 0 && kind != kSubConst;
    int4 na = make_int4(constant, constant, constant, constant), nb = na;
    if (loads) {
        na = *reinterpret_cast<const int4 *>(row0);
        if (!NARROW) nb = *reinterpret_cast<const int4 *>(row0 + half_off);
    }
    const bool folded = NARROW && loads;  // na = eight int16 values (constant subframes keep the value itself)
    uint32_t lossy = 0, buf = 0;
    uint8_t *const slot0 = tile + g * (chunk_bytes + 16) + c * CB;
    const uint32_t rs = C * CB;
    // The loop exists once per predictor class (the class is warp-uniform and fixed for the launch's warp): with one loop
    // and a dispatch inside it, the register window had one layout for all classes and every iteration paid the branches
    // and ~20 moves to and from it (r06l).
    auto run = [&](auto pc_tag) {
    constexpr uint32_t PC = decltype(pc_tag)::value;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < bs_max; i0 += 8) {
        int32_t r[8] = {na.x, na.y, na.z, na.w, nb.x, nb.y, nb.z, nb.w};
        if (folded) {
            const uint32_t w[4] = {(uint32_t)na.x, (uint32_t)na.y, (uint32_t)na.z, (uint32_t)na.w};
#pragma unroll
            for (int u = 0; u < 8; ++u) r[u] = (u & 1) ? ((int32_t)w[u >> 1] >> 16) : (int32_t)(int16_t)w[u >> 1];
        }
        if (loads && i0 + 8 < bs) {  // next group (whole groups are in bounds: the frame's scratch covers roundup8(bs) steps)
            const uint32_t t = (i0 >> 3) + 1;
            const uint8_t *nx = row0 + (size_t)(t >> row_shift) * row_stride + (t & row_mask) * 16u;
            na = *reinterpret_cast<const int4 *>(nx);
            if (!NARROW) nb = *reinterpret_cast<const int4 *>(nx + half_off);
        }
        int32_t y[8];
        if (i0 < order_max) {  // warm-up steps present (first one or two iterations; four in the LARGE kernel)
            if (LARGE) {
                if (PC == 0) lpc_chunk<HMAX, 32, int32_t, true>(h, cf, r, shift, order, i0, y);
                else lpc_chunk<HMAX, 32, int64_t, true>(h, cf, r, shift, order, i0, y);
            } else {
                if (PC <= 2) lpc_chunk<HMAX, 12, int32_t, true>(h, cf, r, shift, order, i0, y);
                else lpc_chunk<HMAX, 12, int64_t, true>(h, cf, r, shift, order, i0, y);
            }
        } else if (LARGE) {
            if (PC == 0) lpc_chunk<HMAX, 32, int32_t, false>(h, cf, r, shift, order, i0, y);
            else lpc_chunk<HMAX, 32, int64_t, false>(h, cf, r, shift, order, i0, y);
        } else {
            if (PC == 0) lpc_chunk<HMAX, 4, int32_t, false>(h, cf, r, shift, order, i0, y);
            else if (PC == 1) lpc_chunk<HMAX, 8, int32_t, false>(h, cf, r, shift, order, i0, y);
            else if (PC == 2) lpc_chunk<HMAX, 12, int32_t, false>(h, cf, r, shift, order, i0, y);
            else if (PC == 3) lpc_chunk<HMAX, 8, int64_t, false>(h, cf, r, shift, order, i0, y);
            else lpc_chunk<HMAX, 12, int64_t, false>(h, cf, r, shift, order, i0, y);
        }

        // ---- wasted bits, decorrelation, narrowing, MSB-align shift
        int32_t o[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) o[u] = y[u];
        if (LARGE || any_wasted) {  // warp-uniform: streams without wasted bits skip the shifts (the 32-tap kernel is faster without the branch)
#pragma unroll
            for (int u = 0; u < 8; ++u) o[u] = narrow<CB>((int32_t)((uint32_t)y[u] << wasted));  // `<<=` on the container type (:494, :537): wraps
        }
        if (CB == 2 && pair16) {
            const bool whole = i0 + 8 <= bs;  // the same for both lanes of a frame
            if (__all_sync(0xFFFFFFFFu, whole || i0 >= bs)) {  // no frame of the warp is in its last, partial group
                int32_t x[4], snd[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    x[j] = odd ? o[j + 4] : o[j];
                    snd[j] = odd ? o[j] : o[j + 4];
                }
                uint32_t w[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int32_t yv = __shfl_xor_sync(0xFFFFFFFFu, snd[j], 1);
                    const int32_t lsb = (odd ? x[j] : yv) & pm;
                    int32_t left = (kxl * x[j] + kyl * yv + lsb) >> psh;
                    int32_t right = (kxr * x[j] + kyr * yv + lsb) >> psh;
                    if (msb_shift != 0) {
                        const int32_t cl = narrow<CB>(left), cr = narrow<CB>(right);
                        left = narrow<CB>((int32_t)((uint32_t)cl << msb_shift));  // `<<=` discards the bits shifted out
                        right = narrow<CB>((int32_t)((uint32_t)cr << msb_shift));
                        if (((left >> msb_shift) != cl || (right >> msb_shift) != cr) && whole) lossy = 1;
                    }
                    w[j] = __byte_perm((uint32_t)left, (uint32_t)right, 0x5410);
                }
                if (whole) *reinterpret_cast<uint4 *>(pair_ptr + (size_t)i0 * 4u) = make_uint4(w[0], w[1], w[2], w[3]);
                rd_ptr[0] += chunk_bytes;  // the staged path's cursor (a later, partial group of some frame uses it)
                continue;
            }
        }
        if (stereo) {  // warp-uniform: every lane takes part in the shuffles
            int32_t p[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) p[u] = __shfl_xor_sync(0xFFFFFFFFu, o[u], 1);
            if (!wide_pair) {
#pragma unroll
                for (int u = 0; u < 8; ++u) o[u] = (da * o[u] + db * p[u] + ((p[u] & mp) | (o[u] & ms))) >> dsh;
            } else {  // 29..31-bit streams: the mid/side sum needs more than 32 bits
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    o[u] = (int32_t)(((int64_t)da * o[u] + (int64_t)db * p[u] + (int64_t)((p[u] & mp) | (o[u] & ms))) >> dsh);
            }
        }
        if (msb_shift != 0) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int32_t cv = narrow<CB>(o[u]);
                const int32_t sh = narrow<CB>((int32_t)((uint32_t)cv << msb_shift));  // `<<=` discards the bits shifted out
                if ((sh >> msb_shift) != cv && i0 + u < bs) lossy = 1;
                o[u] = sh;
            }
        }

        // ---- store: staged PCM byte image -> 16-byte stores, or element by element (tails, odd addresses)
        uint8_t *tb = tile + buf * (kTileBytes / 2);
        if (staged_ok && i0 + 8 <= bs) {
            uint8_t *slot = slot0 + buf * (kTileBytes / 2);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (CB == 2) *reinterpret_cast<int16_t *>(slot + u * rs) = (int16_t)o[u];
                else if (CB == 4) *reinterpret_cast<int32_t *>(slot + u * rs) = o[u];
                else *reinterpret_cast<int8_t *>(slot + u * rs) = (int8_t)o[u];
            }
        } else if (i0 < bs) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (i0 + u < bs) {
                    const uint64_t e = elem_base + (uint64_t)(i0 + u) * C;
                    if (CB == 2) reinterpret_cast<int16_t *>(pcm)[e] = (int16_t)o[u];
                    else if (CB == 4) reinterpret_cast<int32_t *>(pcm)[e] = o[u];
                    else reinterpret_cast<int8_t *>(pcm)[e] = (int8_t)o[u];
                }
            }
        }
        __syncwarp();
        if (i0 + 8 <= rd_bs[0]) {
            if (unit == 16) *reinterpret_cast<uint4 *>(rd_ptr[0]) = *reinterpret_cast<const uint4 *>(tb + rd_soff[0]);
            else *reinterpret_cast<uint2 *>(rd_ptr[0]) = *reinterpret_cast<const uint2 *>(tb + rd_soff[0]);
        }
        rd_ptr[0] += chunk_bytes;
        if (two_pieces) {
            if (i0 + 8 <= rd_bs[1]) {
                if (unit == 16) *reinterpret_cast<uint4 *>(rd_ptr[1]) = *reinterpret_cast<const uint4 *>(tb + rd_soff[1]);
                else *reinterpret_cast<uint2 *>(rd_ptr[1]) = *reinterpret_cast<const uint2 *>(tb + rd_soff[1]);
            }
            rd_ptr[1] += chunk_bytes;
        }
        buf ^= 1u;  // the next iteration's stores go to the other half: one __syncwarp per iteration is enough
    }
    };
    if (LARGE) {
        if (pclass == 0) run(std::integral_constant<uint32_t, 0>{});
        else run(std::integral_constant<uint32_t, 1>{});
    } else {
        if (pclass == 0) run(std::integral_constant<uint32_t, 0>{});
        else if (pclass == 1) run(std::integral_constant<uint32_t, 1>{});
        else if (pclass == 2) run(std::integral_constant<uint32_t, 2>{});
        else if (pclass == 3) run(std::integral_constant<uint32_t, 3>{});
        else run(std::integral_constant<uint32_t, 4>{});
    }
    if (lossy) atomicOr(lossy_flag, 1u);
}

cudaError_t launch_predict_epilogue(const PredictParams &p, cudaStream_t st) {
    if (p.n_frames == 0) return cudaSuccess;
    uint32_t cp_shift = 0;
    while ((1u << cp_shift) < p.max_channels) ++cp_shift;  // lanes per frame: channels rounded up to a power of two
    if (cp_shift > 3) cp_shift = 3;
    const uint32_t G = 32u >> cp_shift;
    const uint32_t blocks = (uint32_t)(((uint64_t)p.n_frames + G - 1) / G);  // one warp per CTA
#define ZF_PRED(LARGE, CB, NARROW)                                                                                               \
    k_predict_epilogue<LARGE, CB, NARROW><<<blocks, 32, 0, st>>>(p.frames, p.outs, p.n_frames, (const int32_t *)p.scratch,      \
                                                                 (const SubRec *)p.subs, (uint8_t *)p.pcm, p.msb_shift, cp_shift, \
                                                                 p.lossy_flag, p.big_orders)
    // narrow scratch: streams of <= 16 bits per sample, i.e. containers of 1 or 2 bytes
    ZF_COUNT_LAUNCH();
    if (p.narrow) {
        if (p.container_bytes == 1) ZF_PRED(false, 1, true);
        else ZF_PRED(false, 2, true);
    } else if (p.container_bytes == 1) ZF_PRED(false, 1, false);
    else if (p.container_bytes == 2) ZF_PRED(false, 2, false);
    else ZF_PRED(false, 4, false);
    ZF_COUNT_LAUNCH();
    if (p.narrow) {
        if (p.container_bytes == 1) ZF_PRED(true, 1, true);
        else ZF_PRED(true, 2, true);
    } else if (p.container_bytes == 1) ZF_PRED(true, 1, false);
    else if (p.container_bytes == 2) ZF_PRED(true, 2, false);
    else ZF_PRED(true, 4, false);
#undef ZF_PRED
    return cudaGetLastError();
}

}  // namespace zf
