// epilogue.cu — K3 epilogue and K4 md5 (sm_100a).
//
// K3 restates, per time step of a frame, the reference's stereo decorrelation
// (src/zflac.zig:553-578), the narrowing store into the container type (:494, :537 — the
// wasted-bit shift was already applied by K2) and the MSB-align shift (:287-306), and writes
// the interleaved, frame-major/channel-minor PCM layout DecodedFLAC returns (:443). It reads
// the planar scratch coalesced and writes whole interleaved time steps per thread.
//
// K4 restates the MD5 gate (:267-280): RFC 1321 over the little-endian container bytes of the
// PRE-shift samples, 3 bytes per sample for 17..24-bit streams (Q4). A single stream's MD5 is
// a serial chain (64 dependent steps per 64-byte block); lanes 1..31 only prefetch and pack.
#include <type_traits>

#include "common.cuh"
#include "kernels.cuh"

namespace zf {

constexpr int kEpiThreads = 128;

// Decorrelate one stereo pair (values are exact integers; int64 covers the 33-bit side channel).
__device__ __forceinline__ void decorrelate(uint32_t ch_code, int64_t &a, int64_t &b) {
    if (ch_code == 8) {         // left/side: R = L - S (:555-560)
        b = a - b;
    } else if (ch_code == 9) {  // side/right: L = S + R (:561-566)
        a = a + b;
    } else if (ch_code == 10) { // mid/side (:567-576)
        const int64_t mid = (a << 1) + (b & 1);
        const int64_t side = b;
        a = (mid + side) >> 1;
        b = (mid - side) >> 1;
    }
}

template <typename CT> __device__ __forceinline__ CT narrow_shift(int64_t v, uint32_t shift, uint32_t &lossy) {
    using UT = typename std::conditional<sizeof(CT) == 1, uint8_t, typename std::conditional<sizeof(CT) == 2, uint16_t, uint32_t>::type>::type;
    const CT c = (CT)(UT)(uint64_t)v;            // @intCast narrowing (wraps where the reference is undefined)
    const CT sh = (CT)(UT)((UT)c << shift);      // `<<=` discards the bits shifted out
    if ((CT)(sh >> shift) != c) lossy = 1;
    return sh;
}

// 4 consecutive samples of one channel plane starting at an index that is a multiple of 4 (planes
// start at multiples of 8 elements, so the vector loads are aligned).
__device__ __forceinline__ void ld_plane4(const int32_t *p, int64_t (&v)[4]) {
    const int4 q = __ldg(reinterpret_cast<const int4 *>(p));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
__device__ __forceinline__ void ld_plane4(const int64_t *p, int64_t (&v)[4]) {
    const longlong2 a = __ldg(reinterpret_cast<const longlong2 *>(p));
    const longlong2 b = __ldg(reinterpret_cast<const longlong2 *>(p + 2));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

// One thread = 4 consecutive time steps x all C channels of one frame: C coalesced 128-bit plane
// loads in, 4*C contiguous container values out (128-bit stores when the PCM address allows).
template <typename ST, typename CT, int C>
__device__ __forceinline__ void epilogue_frame(const FrameRec &fr, const ST *__restrict__ scratch, CT *__restrict__ pcm,
                                               uint32_t shift, uint32_t &lossy) {
    using UT = typename std::conditional<sizeof(CT) == 1, uint8_t, typename std::conditional<sizeof(CT) == 2, uint16_t, uint32_t>::type>::type;
    constexpr int kWords = 4 * C * (int)sizeof(CT) / 4;  // 32-bit words written per thread
    const uint32_t bs = fr.block_size;
    const ST *src = scratch + fr.scratch_base;
    CT *dst = pcm + fr.sample_base;
    for (uint32_t t = 4 * (blockIdx.y * kEpiThreads + threadIdx.x); t < bs; t += 4 * gridDim.y * kEpiThreads) {
        int64_t v[C][4];
#pragma unroll
        for (int c = 0; c < C; ++c) ld_plane4(src + (size_t)c * fr.scratch_pitch + t, v[c]);  // pitch is padded to 8: in bounds
        CT o[4 * C];
        uint32_t ignored = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // the subframe's value went into the container type and took its wasted-bits shift THERE (:494, :537; `<<=` drops
            // what leaves the type): the decorrelation reads it back from the container
#pragma unroll
            for (int c = 0; c < C; ++c) v[c][k] = (int64_t)(CT)(UT)(uint64_t)v[c][k];
            if (C == 2) decorrelate(fr.ch_code, v[0][k], v[C - 1][k]);
#pragma unroll
            for (int c = 0; c < C; ++c) o[k * C + c] = narrow_shift<CT>(v[c][k], shift, t + k < bs ? lossy : ignored);  // plane padding is garbage
        }
        CT *out = dst + (size_t)t * C;
        const uint32_t valid = bs - t < 4 ? bs - t : 4;
        if (valid == 4 && (reinterpret_cast<uintptr_t>(out) & 3) == 0) {
            uint32_t w[kWords];
#pragma unroll
            for (int i = 0; i < kWords; ++i) {
                if (sizeof(CT) == 4) w[i] = (uint32_t)o[i];
                else if (sizeof(CT) == 2) w[i] = (uint32_t)(UT)o[2 * i] | ((uint32_t)(UT)o[2 * i + 1] << 16);
                else w[i] = (uint32_t)(UT)o[4 * i] | ((uint32_t)(UT)o[4 * i + 1] << 8) | ((uint32_t)(UT)o[4 * i + 2] << 16) | ((uint32_t)(UT)o[4 * i + 3] << 24);
            }
            if (kWords % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
#pragma unroll
                for (int i = 0; i < kWords / 4; ++i)
                    reinterpret_cast<uint4 *>(out)[i] = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
            } else if (kWords % 2 == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0) {
#pragma unroll
                for (int i = 0; i < kWords / 2; ++i) reinterpret_cast<uint2 *>(out)[i] = make_uint2(w[2 * i], w[2 * i + 1]);
            } else {
#pragma unroll
                for (int i = 0; i < kWords; ++i) reinterpret_cast<uint32_t *>(out)[i] = w[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4 * C; ++i)
                if (i < (int)valid * C) out[i] = o[i];
        }
    }
}

template <typename ST, typename CT>
__global__ void __launch_bounds__(kEpiThreads)
k_epilogue(const FrameRec *__restrict__ frames, const ST *__restrict__ scratch, CT *__restrict__ pcm, uint32_t shift,
           uint32_t *__restrict__ lossy_flag) {
    const FrameRec fr = frames[blockIdx.x];
    if (fr.block_size == 0) return;  // dropped candidate slot
    uint32_t lossy = 0;
    switch (fr.channels) {
        case 1: epilogue_frame<ST, CT, 1>(fr, scratch, pcm, shift, lossy); break;
        case 2: epilogue_frame<ST, CT, 2>(fr, scratch, pcm, shift, lossy); break;
        case 3: epilogue_frame<ST, CT, 3>(fr, scratch, pcm, shift, lossy); break;
        case 4: epilogue_frame<ST, CT, 4>(fr, scratch, pcm, shift, lossy); break;
        case 5: epilogue_frame<ST, CT, 5>(fr, scratch, pcm, shift, lossy); break;
        case 6: epilogue_frame<ST, CT, 6>(fr, scratch, pcm, shift, lossy); break;
        case 7: epilogue_frame<ST, CT, 7>(fr, scratch, pcm, shift, lossy); break;
        default: epilogue_frame<ST, CT, 8>(fr, scratch, pcm, shift, lossy); break;
    }
    if (lossy) atomicOr(lossy_flag, 1u);
}

template <typename CT>
__global__ void k_shift_inplace(CT *__restrict__ pcm, uint64_t n, uint32_t shift) {
    using UT = typename std::make_unsigned<CT>::type;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        pcm[i] = (CT)(UT)((UT)pcm[i] << shift);
}

cudaError_t launch_epilogue(const EpilogueParams &p, cudaStream_t st) {
    if (p.n_frames == 0) return cudaSuccess;
    uint32_t gy = (p.max_block_size + 4 * kEpiThreads - 1) / (4 * kEpiThreads);
    if (gy > 64) gy = 64;
    if (gy == 0) gy = 1;
    dim3 grid(p.n_frames, gy);
    ZF_COUNT_LAUNCH();
#define ZF_EPI(ST, CT) \
    k_epilogue<ST, CT><<<grid, kEpiThreads, 0, st>>>(p.frames, (const ST *)p.scratch, (CT *)p.pcm, p.msb_shift, p.lossy_flag)
    if (p.wide_scratch) {
        if (p.container_bytes == 1) ZF_EPI(int64_t, int8_t);
        else if (p.container_bytes == 2) ZF_EPI(int64_t, int16_t);
        else ZF_EPI(int64_t, int32_t);
    } else {
        if (p.container_bytes == 1) ZF_EPI(int32_t, int8_t);
        else if (p.container_bytes == 2) ZF_EPI(int32_t, int16_t);
        else ZF_EPI(int32_t, int32_t);
    }
#undef ZF_EPI
    return cudaGetLastError();
}

cudaError_t launch_shift_inplace(void *pcm, uint64_t n_samples, uint32_t container_bytes, uint32_t shift, cudaStream_t st) {
    if (n_samples == 0 || shift == 0) return cudaSuccess;
    const int threads = 256;
    uint64_t want = (n_samples + threads - 1) / threads;
    const uint32_t blocks = (uint32_t)(want < 148ull * 16 ? want : 148ull * 16);
    ZF_COUNT_LAUNCH();
    if (container_bytes == 2) k_shift_inplace<int16_t><<<blocks, threads, 0, st>>>((int16_t *)pcm, n_samples, shift);
    else if (container_bytes == 4) k_shift_inplace<int32_t><<<blocks, threads, 0, st>>>((int32_t *)pcm, n_samples, shift);
    else k_shift_inplace<int8_t><<<blocks, threads, 0, st>>>((int8_t *)pcm, n_samples, shift);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ K4: MD5 (RFC 1321)
__constant__ uint32_t c_md5_k[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
    0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
    0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
    0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
    0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};

#define MD5_STEP(f, a, b, c, d, m, k, s)                     \
    do {                                                     \
        a += f(b, c, d) + (m) + (k);                         \
        a = __funnelshift_l(a, a, s) + b;                    \
    } while (0)
#define MD5_F(x, y, z) ((z) ^ ((x) & ((y) ^ (z))))
#define MD5_G(x, y, z) ((y) ^ ((z) & ((x) ^ (y))))
#define MD5_H(x, y, z) ((x) ^ (y) ^ (z))
#define MD5_I(x, y, z) ((y) ^ ((x) | ~(z)))

__device__ __forceinline__ void md5_block(uint32_t (&st)[4], const uint32_t *__restrict__ M) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    uint32_t m[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) m[i] = M[i];
#define R1(i, a, b, c, d, s) MD5_STEP(MD5_F, a, b, c, d, m[(i)], c_md5_k[(i)], s)
#define R2(i, a, b, c, d, s) MD5_STEP(MD5_G, a, b, c, d, m[(5 * (i) + 1) & 15], c_md5_k[(i)], s)
#define R3(i, a, b, c, d, s) MD5_STEP(MD5_H, a, b, c, d, m[(3 * (i) + 5) & 15], c_md5_k[(i)], s)
#define R4(i, a, b, c, d, s) MD5_STEP(MD5_I, a, b, c, d, m[(7 * (i)) & 15], c_md5_k[(i)], s)
#pragma unroll
    for (int i = 0; i < 16; i += 4) {
        R1(i + 0, a, b, c, d, 7); R1(i + 1, d, a, b, c, 12); R1(i + 2, c, d, a, b, 17); R1(i + 3, b, c, d, a, 22);
    }
#pragma unroll
    for (int i = 16; i < 32; i += 4) {
        R2(i + 0, a, b, c, d, 5); R2(i + 1, d, a, b, c, 9); R2(i + 2, c, d, a, b, 14); R2(i + 3, b, c, d, a, 20);
    }
#pragma unroll
    for (int i = 32; i < 48; i += 4) {
        R3(i + 0, a, b, c, d, 4); R3(i + 1, d, a, b, c, 11); R3(i + 2, c, d, a, b, 16); R3(i + 3, b, c, d, a, 23);
    }
#pragma unroll
    for (int i = 48; i < 64; i += 4) {
        R4(i + 0, a, b, c, d, 6); R4(i + 1, d, a, b, c, 10); R4(i + 2, c, d, a, b, 15); R4(i + 3, b, c, d, a, 21);
    }
#undef R1
#undef R2
#undef R3
#undef R4
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}

// Byte `m` of the padded MD5 message of a job.
template <int HB> __device__ __forceinline__ uint32_t md5_msg_byte(const Md5Job &j, uint64_t m, uint64_t msg_len, uint64_t padded_len) {
    if (m < msg_len) {
        const uint64_t smp = m / HB;
        const uint32_t byt = (uint32_t)(m % HB);
        int32_t v;
        if (j.container_bytes == 1) v = ((const int8_t *)j.pcm)[smp];
        else if (j.container_bytes == 2) v = ((const int16_t *)j.pcm)[smp];
        else v = ((const int32_t *)j.pcm)[smp];
        v >>= j.unshift;
        return ((uint32_t)v >> (8 * byt)) & 0xFFu;
    }
    if (m == msg_len) return 0x80u;
    if (m >= padded_len - 8) return (uint32_t)(((msg_len * 8) >> (8 * (m - (padded_len - 8)))) & 0xFFu);
    return 0u;
}

constexpr int kMd5ChunkBlocks = 32;  // 64-byte blocks staged per round: one per lane

template <int HB>
__device__ void md5_run(const Md5Job &j, uint8_t *digest, uint32_t (*stage)[16]) {
    const int lane = threadIdx.x & 31;
    const uint64_t msg_len = j.n_samples * HB;
    const uint64_t padded_len = (msg_len + 9 + 63) & ~63ull;
    const uint64_t n_blocks = padded_len >> 6;
    const bool raw = (HB == (int)j.container_bytes) && j.unshift == 0 && ((uintptr_t)j.pcm & 3) == 0;
    uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    for (uint64_t b0 = 0; b0 < n_blocks; b0 += kMd5ChunkBlocks) {
        const uint64_t blk = b0 + lane;
        if (blk < n_blocks) {
            const uint64_t byte0 = blk << 6;
            if (raw && byte0 + 64 <= msg_len) {  // the message is the PCM memory itself
                const uint4 *p = reinterpret_cast<const uint4 *>((const uint8_t *)j.pcm + byte0);
                if (((uintptr_t)p & 15) == 0) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint4 v = __ldg(p + q);
                        stage[lane][4 * q + 0] = v.x; stage[lane][4 * q + 1] = v.y;
                        stage[lane][4 * q + 2] = v.z; stage[lane][4 * q + 3] = v.w;
                    }
                } else {
                    const uint32_t *pw = reinterpret_cast<const uint32_t *>((const uint8_t *)j.pcm + byte0);
#pragma unroll
                    for (int q = 0; q < 16; ++q) stage[lane][q] = __ldg(pw + q);
                }
            } else {
                for (int wi = 0; wi < 16; ++wi) {
                    uint32_t w = 0;
#pragma unroll
                    for (int bb = 0; bb < 4; ++bb) w |= md5_msg_byte<HB>(j, byte0 + 4 * wi + bb, msg_len, padded_len) << (8 * bb);
                    stage[lane][wi] = w;
                }
            }
        }
        __syncwarp();
        if (lane == 0) {
            const int nb = (n_blocks - b0) < (uint64_t)kMd5ChunkBlocks ? (int)(n_blocks - b0) : kMd5ChunkBlocks;
            for (int i = 0; i < nb; ++i) md5_block(st, stage[i]);
        }
        __syncwarp();
    }
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int k = 0; k < 4; ++k) digest[4 * i + k] = (uint8_t)(st[i] >> (8 * k));
    }
}

// Batch variant: one LANE per stream. Every chain is serial, but 32 of them share each instruction; with thousands
// of short files this is ~30x fewer issue slots than one warp per chain (lanes 1..31 idle during md5_block).
//
// The message bytes reach the lanes through shared memory: per round the warp copies the next 512 bytes of EVERY lane's
// stream with cp.async — for each stream one instruction whose 32 lanes read 512 contiguous bytes — one round ahead of
// the hashing, and a lane then reads its own row. Loading from the lanes themselves (each lane four 16-byte loads from
// its own file per block, r02z) cost 2 350 cycles per block where the chain needs 1 055: every load instruction touched
// 32 different pages, and its slowest lane set the pace.
constexpr int kLaneChunk = 512;            // bytes per stream and round: eight 64-byte blocks
constexpr int kLaneRow = kLaneChunk + 16;  // row pitch: the 16-byte loads of eight neighbouring lanes hit different banks

// the blocks from `first_block` on (what is not PCM memory as it is: repacked samples, the padding), one lane's own way
template <int HB>
__device__ void md5_lane_tail(const Md5Job &j, uint32_t (&st)[4], uint64_t first_block) {
    const uint64_t msg_len = j.n_samples * HB;
    const uint64_t padded_len = (msg_len + 9 + 63) & ~63ull;
    const uint64_t n_blocks = padded_len >> 6;
    uint32_t m[16];
    for (uint64_t blk = first_block; blk < n_blocks; ++blk) {
        const uint64_t byte0 = blk << 6;
        uint32_t slow[16];  // dynamic indexing stays in this (local-memory) array: m[] keeps its registers
        for (int wi = 0; wi < 16; ++wi) {
            uint32_t w = 0;
#pragma unroll
            for (int bb = 0; bb < 4; ++bb) w |= md5_msg_byte<HB>(j, byte0 + 4 * wi + bb, msg_len, padded_len) << (8 * bb);
            slow[wi] = w;
        }
#pragma unroll
        for (int wi = 0; wi < 16; ++wi) m[wi] = slow[wi];
        md5_block(st, m);
    }
}

__global__ void __launch_bounds__(32)
k_md5_lanes(const Md5Job *__restrict__ jobs, uint32_t n_jobs, uint8_t *__restrict__ digests) {
    __shared__ __align__(16) uint8_t stage[2][32][kLaneRow];
    const uint32_t lane = threadIdx.x;
    const uint32_t job = blockIdx.x * 32 + lane;
    Md5Job j;
    j.pcm = nullptr;
    j.n_samples = 0;
    j.container_bytes = j.hash_bytes = 2;
    j.unshift = j.pad = 0;
    if (job < n_jobs) j = jobs[job];
    const uint32_t hb = j.hash_bytes;
    // whole blocks that are PCM memory as it is
    const bool raw = job < n_jobs && hb == j.container_bytes && j.unshift == 0 && ((uintptr_t)j.pcm & 15) == 0;
    const uint64_t raw_blocks = raw ? (j.n_samples * hb) >> 6 : 0;
    const uint64_t raw_bytes = raw_blocks << 6;
    const uint32_t my_rounds = (uint32_t)((raw_blocks + 7) >> 3);  // < 2^32: a stream of 2 TB would not be in a batch
    const uint32_t n_rounds = __reduce_max_sync(0xFFFFFFFFu, my_rounds);
    const unsigned long long my_ptr = (unsigned long long)(uintptr_t)j.pcm;
    uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};

    auto issue = [&](uint32_t round, int buf) {  // 512 bytes of every stream, 16 per lane, zero-filled past a stream's raw part
        const unsigned long long off = (unsigned long long)round * kLaneChunk + lane * 16u;
#pragma unroll 4
        for (int o = 0; o < 32; ++o) {
            const unsigned long long po = __shfl_sync(0xFFFFFFFFu, my_ptr, o);
            const unsigned long long rb = __shfl_sync(0xFFFFFFFFu, (unsigned long long)raw_bytes, o);
            const uint32_t n = off + 16u <= rb ? 16u : 0u;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&stage[buf][o][lane * 16u]);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(po + (n ? off : 0ull)), "r"(n) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (n_rounds) issue(0, 0);
    for (uint32_t r = 0; r < n_rounds; ++r) {
        if (r + 1 < n_rounds) {  // the next round is in flight while this one is hashed
            issue(r + 1, (r + 1) & 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        const uint4 *row = reinterpret_cast<const uint4 *>(stage[r & 1][lane]);
#pragma unroll 1
        for (int b = 0; b < 8; ++b) {
            if ((uint64_t)r * 8 + b < raw_blocks) {
                uint32_t m[16];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint4 v = row[4 * b + q];
                    m[4 * q + 0] = v.x; m[4 * q + 1] = v.y; m[4 * q + 2] = v.z; m[4 * q + 3] = v.w;
                }
                md5_block(st, m);
            }
        }
        __syncwarp();  // everyone is done with this buffer before round r + 2 lands in it
    }
    if (job >= n_jobs) return;
    switch (hb) {
        case 1: md5_lane_tail<1>(j, st, raw_blocks); break;
        case 2: md5_lane_tail<2>(j, st, raw_blocks); break;
        case 3: md5_lane_tail<3>(j, st, raw_blocks); break;
        default: md5_lane_tail<4>(j, st, raw_blocks); break;
    }
    uint8_t *digest = digests + 16 * (size_t)job;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int k = 0; k < 4; ++k) digest[4 * i + k] = (uint8_t)(st[i] >> (8 * k));
}

constexpr int kMd5WarpsPerBlock = 2;

__global__ void __launch_bounds__(32 * kMd5WarpsPerBlock)
k_md5(const Md5Job *__restrict__ jobs, uint32_t n_jobs, uint8_t *__restrict__ digests) {
    __shared__ uint32_t stage[kMd5WarpsPerBlock][kMd5ChunkBlocks][16];
    const uint32_t job = blockIdx.x * kMd5WarpsPerBlock + (threadIdx.x >> 5);
    if (job >= n_jobs) return;
    const Md5Job j = jobs[job];
    uint32_t(*stg)[16] = stage[threadIdx.x >> 5];
    uint8_t *dg = digests + 16 * (size_t)job;
    switch (j.hash_bytes) {
        case 1: md5_run<1>(j, dg, stg); break;
        case 2: md5_run<2>(j, dg, stg); break;
        case 3: md5_run<3>(j, dg, stg); break;
        default: md5_run<4>(j, dg, stg); break;
    }
}

cudaError_t launch_md5(const Md5Job *d_jobs, uint32_t n_jobs, uint8_t *d_digests, cudaStream_t st) {
    if (n_jobs == 0) return cudaSuccess;
    ZF_COUNT_LAUNCH();
    if (n_jobs >= 256) {  // enough chains to give every lane its own
        k_md5_lanes<<<(n_jobs + 31) / 32, 32, 0, st>>>(d_jobs, n_jobs, d_digests);
        return cudaGetLastError();
    }
    k_md5<<<(n_jobs + kMd5WarpsPerBlock - 1) / kMd5WarpsPerBlock, 32 * kMd5WarpsPerBlock, 0, st>>>(d_jobs, n_jobs, d_digests);
    return cudaGetLastError();
}

}  // namespace zf
