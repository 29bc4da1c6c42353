// hostmd5.h — the one host-side VERIFICATION stage of the library: the STREAMINFO MD5 gate of a single long stream
// (reference src/zflac.zig:267-280, RFC 1321) computed on a host thread from the PCM bytes as they land in host memory.
//
// Why on the host: one stream's MD5 is a Merkle-Damgard chain, 64 dependent steps of 4 dependent ALU ops per 64-byte
// block, and nothing in it can run in parallel. A GPU thread pays ~4 cycles per dependent op at 1.9 GHz: 1 055 cycles
// per block measured = 119 MB/s whatever the number of SMs (k_md5, epilogue.cu); a host core pays one cycle per op at
// 3+ GHz: 550-700 MB/s. So for ONE long stream the chain runs where a dependent op is cheapest, overlapped with the
// device->host copy it reads from. This is not a decode fallback: no sample is ever produced on the CPU, the hash only
// READS the PCM the kernels wrote, and a batch of many streams still hashes on the device (one chain per lane,
// k_md5_lanes). ZFLACB200_MD5_ON_DEVICE forces the device chain for single streams too.
#pragma once
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>

#include <cuda_runtime.h>

namespace zf {

struct Md5 {
    uint32_t a = 0x67452301u, b = 0xefcdab89u, c = 0x98badcfeu, d = 0x10325476u;
    uint64_t n_bytes = 0;
    uint8_t buf[64];
    uint32_t fill = 0;

    static inline uint32_t rol(uint32_t v, int s) { return (v << s) | (v >> (32 - s)); }

    // `blocks` 64-byte blocks from p (any alignment)
    void compress(const uint8_t *p, size_t blocks) {
        uint32_t A = a, B = b, C = c, D = d;
        for (; blocks; --blocks, p += 64) {
            uint32_t m[16];
            memcpy(m, p, 64);  // little-endian host
            const uint32_t a0 = A, b0 = B, c0 = C, d0 = D;
            // F(b,c,d) = d ^ (b & (c ^ d)): (c ^ d) is off the chain through b
#define ZF_MD5_F(a, b, c, d, k, s, t) a += (d ^ (b & (c ^ d))) + m[k] + t; a = rol(a, s) + b;
            // G(b,c,d) = (d & b) | (~d & c): the ~d & c term is off the chain, and the two terms have no common bit -> add
#define ZF_MD5_G(a, b, c, d, k, s, t) a += (~d & c) + m[k] + t; a += (d & b); a = rol(a, s) + b;
#define ZF_MD5_H(a, b, c, d, k, s, t) a += (b ^ c ^ d) + m[k] + t; a = rol(a, s) + b;
#define ZF_MD5_I(a, b, c, d, k, s, t) a += (c ^ (b | ~d)) + m[k] + t; a = rol(a, s) + b;
            ZF_MD5_F(A, B, C, D, 0, 7, 0xd76aa478u) ZF_MD5_F(D, A, B, C, 1, 12, 0xe8c7b756u) ZF_MD5_F(C, D, A, B, 2, 17, 0x242070dbu)
            ZF_MD5_F(B, C, D, A, 3, 22, 0xc1bdceeeu) ZF_MD5_F(A, B, C, D, 4, 7, 0xf57c0fafu) ZF_MD5_F(D, A, B, C, 5, 12, 0x4787c62au)
            ZF_MD5_F(C, D, A, B, 6, 17, 0xa8304613u) ZF_MD5_F(B, C, D, A, 7, 22, 0xfd469501u) ZF_MD5_F(A, B, C, D, 8, 7, 0x698098d8u)
            ZF_MD5_F(D, A, B, C, 9, 12, 0x8b44f7afu) ZF_MD5_F(C, D, A, B, 10, 17, 0xffff5bb1u) ZF_MD5_F(B, C, D, A, 11, 22, 0x895cd7beu)
            ZF_MD5_F(A, B, C, D, 12, 7, 0x6b901122u) ZF_MD5_F(D, A, B, C, 13, 12, 0xfd987193u) ZF_MD5_F(C, D, A, B, 14, 17, 0xa679438eu)
            ZF_MD5_F(B, C, D, A, 15, 22, 0x49b40821u)
            ZF_MD5_G(A, B, C, D, 1, 5, 0xf61e2562u) ZF_MD5_G(D, A, B, C, 6, 9, 0xc040b340u) ZF_MD5_G(C, D, A, B, 11, 14, 0x265e5a51u)
            ZF_MD5_G(B, C, D, A, 0, 20, 0xe9b6c7aau) ZF_MD5_G(A, B, C, D, 5, 5, 0xd62f105du) ZF_MD5_G(D, A, B, C, 10, 9, 0x02441453u)
            ZF_MD5_G(C, D, A, B, 15, 14, 0xd8a1e681u) ZF_MD5_G(B, C, D, A, 4, 20, 0xe7d3fbc8u) ZF_MD5_G(A, B, C, D, 9, 5, 0x21e1cde6u)
            ZF_MD5_G(D, A, B, C, 14, 9, 0xc33707d6u) ZF_MD5_G(C, D, A, B, 3, 14, 0xf4d50d87u) ZF_MD5_G(B, C, D, A, 8, 20, 0x455a14edu)
            ZF_MD5_G(A, B, C, D, 13, 5, 0xa9e3e905u) ZF_MD5_G(D, A, B, C, 2, 9, 0xfcefa3f8u) ZF_MD5_G(C, D, A, B, 7, 14, 0x676f02d9u)
            ZF_MD5_G(B, C, D, A, 12, 20, 0x8d2a4c8au)
            ZF_MD5_H(A, B, C, D, 5, 4, 0xfffa3942u) ZF_MD5_H(D, A, B, C, 8, 11, 0x8771f681u) ZF_MD5_H(C, D, A, B, 11, 16, 0x6d9d6122u)
            ZF_MD5_H(B, C, D, A, 14, 23, 0xfde5380cu) ZF_MD5_H(A, B, C, D, 1, 4, 0xa4beea44u) ZF_MD5_H(D, A, B, C, 4, 11, 0x4bdecfa9u)
            ZF_MD5_H(C, D, A, B, 7, 16, 0xf6bb4b60u) ZF_MD5_H(B, C, D, A, 10, 23, 0xbebfbc70u) ZF_MD5_H(A, B, C, D, 13, 4, 0x289b7ec6u)
            ZF_MD5_H(D, A, B, C, 0, 11, 0xeaa127fau) ZF_MD5_H(C, D, A, B, 3, 16, 0xd4ef3085u) ZF_MD5_H(B, C, D, A, 6, 23, 0x04881d05u)
            ZF_MD5_H(A, B, C, D, 9, 4, 0xd9d4d039u) ZF_MD5_H(D, A, B, C, 12, 11, 0xe6db99e5u) ZF_MD5_H(C, D, A, B, 15, 16, 0x1fa27cf8u)
            ZF_MD5_H(B, C, D, A, 2, 23, 0xc4ac5665u)
            ZF_MD5_I(A, B, C, D, 0, 6, 0xf4292244u) ZF_MD5_I(D, A, B, C, 7, 10, 0x432aff97u) ZF_MD5_I(C, D, A, B, 14, 15, 0xab9423a7u)
            ZF_MD5_I(B, C, D, A, 5, 21, 0xfc93a039u) ZF_MD5_I(A, B, C, D, 12, 6, 0x655b59c3u) ZF_MD5_I(D, A, B, C, 3, 10, 0x8f0ccc92u)
            ZF_MD5_I(C, D, A, B, 10, 15, 0xffeff47du) ZF_MD5_I(B, C, D, A, 1, 21, 0x85845dd1u) ZF_MD5_I(A, B, C, D, 8, 6, 0x6fa87e4fu)
            ZF_MD5_I(D, A, B, C, 15, 10, 0xfe2ce6e0u) ZF_MD5_I(C, D, A, B, 6, 15, 0xa3014314u) ZF_MD5_I(B, C, D, A, 13, 21, 0x4e0811a1u)
            ZF_MD5_I(A, B, C, D, 4, 6, 0xf7537e82u) ZF_MD5_I(D, A, B, C, 11, 10, 0xbd3af235u) ZF_MD5_I(C, D, A, B, 2, 15, 0x2ad7d2bbu)
            ZF_MD5_I(B, C, D, A, 9, 21, 0xeb86d391u)
#undef ZF_MD5_F
#undef ZF_MD5_G
#undef ZF_MD5_H
#undef ZF_MD5_I
            A += a0; B += b0; C += c0; D += d0;
        }
        a = A; b = B; c = C; d = D;
    }

    void update(const uint8_t *p, size_t n) {
        n_bytes += n;
        if (fill) {
            const size_t take = n < 64 - fill ? n : 64 - fill;
            memcpy(buf + fill, p, take);
            fill += (uint32_t)take;
            p += take;
            n -= take;
            if (fill < 64) return;
            compress(buf, 1);
            fill = 0;
        }
        const size_t blocks = n / 64;
        if (blocks) compress(p, blocks);
        p += blocks * 64;
        n -= blocks * 64;
        if (n) {
            memcpy(buf, p, n);
            fill = (uint32_t)n;
        }
    }

    void finish(uint8_t out[16]) {
        const uint64_t bits = n_bytes * 8;
        uint8_t pad[72] = {0x80};
        const size_t padlen = (fill < 56 ? 56 : 120) - fill;
        memcpy(pad + padlen, &bits, 8);  // little-endian host
        const uint64_t keep = n_bytes;
        update(pad, padlen + 8);
        n_bytes = keep;
        memcpy(out, &a, 4);
        memcpy(out + 4, &b, 4);
        memcpy(out + 8, &c, 4);
        memcpy(out + 12, &d, 4);
    }
};

// The reference's packing rule (:268-277, Q4): of each container-typed sample (after undoing the MSB-align shift) the low
// `hash_bytes` bytes enter the hash, little-endian. container == hash_bytes and no shift: the buffer is hashed as it is.
struct PcmHasher {
    Md5 md5;
    uint32_t container_bytes = 2, hash_bytes = 2, unshift = 0;

    void update(const void *pcm, uint64_t n_samples) {
        const uint8_t *p = (const uint8_t *)pcm;
        if (container_bytes == hash_bytes && unshift == 0) {
            md5.update(p, (size_t)(n_samples * container_bytes));
            return;
        }
        uint8_t tmp[3072];  // a multiple of 1, 2, 3 and 4
        const uint32_t per = (uint32_t)(sizeof(tmp) / hash_bytes);
        while (n_samples) {
            const uint32_t m = (uint32_t)(n_samples < per ? n_samples : per);
            uint8_t *q = tmp;
            if (container_bytes == 4) {
                const int32_t *s = (const int32_t *)p;
                if (hash_bytes == 3) {
                    for (uint32_t i = 0; i < m; ++i, q += 3) {
                        const uint32_t v = (uint32_t)(s[i] >> unshift);
                        q[0] = (uint8_t)v; q[1] = (uint8_t)(v >> 8); q[2] = (uint8_t)(v >> 16);
                    }
                } else {
                    for (uint32_t i = 0; i < m; ++i, q += 4) {
                        const int32_t v = s[i] >> unshift;
                        memcpy(q, &v, 4);
                    }
                }
            } else if (container_bytes == 2) {
                const int16_t *s = (const int16_t *)p;
                for (uint32_t i = 0; i < m; ++i, q += 2) {
                    const int16_t v = (int16_t)(s[i] >> unshift);
                    memcpy(q, &v, 2);
                }
            } else {
                const int8_t *s = (const int8_t *)p;
                for (uint32_t i = 0; i < m; ++i) *q++ = (uint8_t)(s[i] >> unshift);
            }
            md5.update(tmp, (size_t)(q - tmp));
            p += (size_t)m * container_bytes;
            n_samples -= m;
        }
    }
    void finish(uint8_t out[16]) { md5.finish(out); }
};

// Hashes pieces of a host buffer in order on its own thread; a piece is read once the CUDA event that marks the end of
// its device->host copy has fired. push() never blocks; finish() joins and yields the digest.
class HashWorker {
  public:
    HashWorker(int device, uint32_t container_bytes, uint32_t hash_bytes, uint32_t unshift) : device_(device) {
        hasher_.container_bytes = container_bytes;
        hasher_.hash_bytes = hash_bytes;
        hasher_.unshift = unshift;
        thread_ = std::thread([this] { run(); });
    }
    ~HashWorker() {
        if (thread_.joinable()) {
            abort_ = true;
            push(nullptr, nullptr, 0, true);
            thread_.join();
        }
    }
    void push(cudaEvent_t after, const void *p, uint64_t n_samples, bool last = false) {
        {
            std::lock_guard<std::mutex> lk(m_);
            q_.push_back(Job{after, p, n_samples, last});
        }
        cv_.notify_one();
    }
    // false: an event wait failed (the digest is then meaningless)
    bool finish(uint8_t out[16]) {
        push(nullptr, nullptr, 0, true);
        thread_.join();
        hasher_.finish(out);
        return ok_;
    }

  private:
    struct Job {
        cudaEvent_t after;
        const void *p;
        uint64_t n;
        bool last;
    };
    void run() {
        cudaSetDevice(device_);
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [this] { return !q_.empty(); });
                j = q_.front();
                q_.pop_front();
            }
            if (j.last) return;
            if (abort_) continue;
            if (j.after && cudaEventSynchronize(j.after) != cudaSuccess) {
                ok_ = false;
                continue;
            }
            if (ok_) hasher_.update(j.p, j.n);
        }
    }
    int device_;
    PcmHasher hasher_;
    std::thread thread_;
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<Job> q_;
    bool ok_ = true;
    volatile bool abort_ = false;
};

}  // namespace zf
