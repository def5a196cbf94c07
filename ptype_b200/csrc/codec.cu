// codec.cu — the data formats either side of the dispatch path, on the device:
//   * "Receiver.Method" strings -> (method, recv) hashes, the way net/rpc's server splits
//     them (strings.LastIndex(".") in readRequestHeader [Go 1.13 net/rpc]) — SURVEY.md N2;
//   * reply values -> encoding/gob messages (`LL 04 00 <int>`), and gob Args bodies ->
//     int64 triples [Go encoding/gob; SURVEY.md appendix A] — SURVEY.md N1.
// Byte/integer work, HBM-bound; no tensor cores.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ptype_b200.h"
#include "device_image.h"

namespace ptype {
namespace {

constexpr int kCodecThreads = 256;

// ------------------------------------------------------------ name resolution
// One thread per fixed-stride, NUL-padded name.  FNV-1a-32 is a serial chain, so a thread
// streams its own string 16 bytes at a time; the hash of the text before the LAST '.' is
// the running hash remembered each time a '.' goes by.
__global__ void __launch_bounds__(kCodecThreads)
    resolve_names_kernel(const unsigned char* __restrict__ names, const uint32_t stride, const size_t n,
                         ptype_method_ref* __restrict__ out, ptype_req* __restrict__ patch) {
    const size_t j = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const unsigned char* s = names + j * stride;
    uint32_t h = 0x811c9dc5u, h_at_dot = 0;
    bool dot = false, done = false;
    for (uint32_t off = 0; off < stride && !done; off += 16) {
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(s + off));  // stride is a multiple of 16
        const uint32_t words[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const uint32_t b = (words[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
            if (done || b == 0) {
                done = true;
            } else {
                if (b == '.') {
                    dot = true;
                    h_at_dot = h;
                }
                h = (h ^ b) * 0x01000193u;
            }
        }
    }
    ptype_method_ref r;
    r.method = h;
    r.recv = dot ? h_at_dot : PTYPE_RECV_ILL_FORMED;
    if (out) out[j] = r;
    if (patch) {
        patch[j].method = r.method;
        patch[j].recv = r.recv;
    }
}

// ------------------------------------------------------------------ gob encode
// encoding/gob signed int: u = i<<1 (i >= 0) or (^i<<1)|1; u < 128 -> one byte, else a byte
// holding the negated byte count followed by the big-endian bytes.
__device__ __forceinline__ uint64_t gob_zigzag(int64_t i) {
    return i < 0 ? ((~static_cast<uint64_t>(i)) << 1) | 1ull : static_cast<uint64_t>(i) << 1;
}
__device__ __forceinline__ uint32_t gob_uint_len(uint64_t u) {
    if (u < 128) return 1;
    return 1 + ((64 - __clzll(static_cast<long long>(u))) + 7) / 8;
}

// pass 1: message length per reply (0 for replies whose status is not OK) + per-block sums
__global__ void __launch_bounds__(kCodecThreads)
    gob_reply_lengths_kernel(const uint4* __restrict__ replies, const size_t n, uint32_t* __restrict__ lens,
                             unsigned long long* __restrict__ block_sums) {
    const size_t j = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    uint32_t len = 0;
    if (j < n) {
        const uint4 r = __ldg(replies + j);
        if (r.w == PTYPE_ST_OK) {
            const int64_t v = static_cast<int64_t>(static_cast<uint64_t>(r.x) | (static_cast<uint64_t>(r.y) << 32));
            len = 3 + gob_uint_len(gob_zigzag(v));  // LL 04 00 <int>
        }
        lens[j] = len;
    }
    uint32_t s = len;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    __shared__ uint32_t part[kCodecThreads / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = 0;
        for (int w = 0; w < kCodecThreads / 32; ++w) b += part[w];
        block_sums[blockIdx.x] = b;
    }
}

// pass 2: exclusive scan of the block sums by one CTA (n_blocks is n/256: a few thousand)
__global__ void __launch_bounds__(1024)
    scan_block_sums_kernel(unsigned long long* __restrict__ block_sums, const size_t n_blocks,
                           unsigned long long* __restrict__ total) {
    __shared__ unsigned long long carry;
    __shared__ unsigned long long warp_tot[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (size_t base = 0; base < n_blocks; base += 1024) {
        const size_t i = base + threadIdx.x;
        const unsigned long long v = i < n_blocks ? block_sums[i] : 0ull;
        unsigned long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if ((threadIdx.x & 31) >= static_cast<unsigned>(o)) incl += t;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
        __syncthreads();
        unsigned long long before = carry;
        for (unsigned w = 0; w < (threadIdx.x >> 5); ++w) before += warp_tot[w];
        if (i < n_blocks) block_sums[i] = before + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

// pass 3: offsets within the block, then the bytes
__global__ void __launch_bounds__(kCodecThreads)
    gob_reply_write_kernel(const uint4* __restrict__ replies, const size_t n, const uint32_t* __restrict__ lens,
                           const unsigned long long* __restrict__ block_offsets,
                           unsigned long long* __restrict__ offsets, uint8_t* __restrict__ out,
                           const unsigned long long out_cap) {
    const size_t j = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint32_t len = j < n ? lens[j] : 0;
    uint32_t incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if ((threadIdx.x & 31) >= static_cast<unsigned>(o)) incl += t;
    }
    __shared__ uint32_t warp_tot[kCodecThreads / 32];
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    unsigned long long at = block_offsets[blockIdx.x] + incl - len;
    for (unsigned w = 0; w < (threadIdx.x >> 5); ++w) at += warp_tot[w];
    if (j >= n) return;
    offsets[j] = at;
    if (j == n - 1) offsets[n] = at + len;
    if (len == 0 || at + len > out_cap) return;
    const uint4 r = __ldg(replies + j);
    const uint64_t u = gob_zigzag(static_cast<int64_t>(static_cast<uint64_t>(r.x) | (static_cast<uint64_t>(r.y) << 32)));
    uint8_t* p = out + at;
    p[0] = static_cast<uint8_t>(len - 1);  // message length (always < 128 here)
    p[1] = 0x04;                           // type id 2 (int) as a gob int
    p[2] = 0x00;                           // top-level non-struct marker
    if (u < 128) {
        p[3] = static_cast<uint8_t>(u);
    } else {
        const uint32_t nb = len - 4;
        p[3] = static_cast<uint8_t>(0u - nb);
        for (uint32_t k = 0; k < nb; ++k) p[4 + k] = static_cast<uint8_t>(u >> (8 * (nb - 1 - k)));
    }
}

// ------------------------------------------------------------------ gob decode
// Args value bodies: (field delta, gob int)* 00, zero-valued fields omitted.  offsets[j] is
// where body j starts; one thread per body.  status 0 = ok, 1 = malformed.
__device__ __forceinline__ bool gob_read_uint(const uint8_t* p, const uint8_t* end, uint32_t& pos, uint64_t& v) {
    if (p + pos >= end) return false;
    const uint8_t b = p[pos++];
    if (b < 128) {
        v = b;
        return true;
    }
    const uint32_t nb = static_cast<uint8_t>(0u - b);
    if (nb < 1 || nb > 8 || p + pos + nb > end) return false;
    uint64_t x = 0;
    for (uint32_t k = 0; k < nb; ++k) x = (x << 8) | p[pos++];
    v = x;
    return true;
}

__global__ void __launch_bounds__(kCodecThreads)
    gob_args_decode_kernel(const uint8_t* __restrict__ bytes, const unsigned long long* __restrict__ offsets,
                           const size_t n, int64_t* __restrict__ args, const int args_per_call,
                           uint8_t* __restrict__ status) {
    const size_t j = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint8_t* p = bytes + offsets[j];
    const uint8_t* end = bytes + offsets[j + 1];
    int64_t a[3] = {0, 0, 0};
    int field = -1;
    uint32_t pos = 0;
    bool ok = true;
    for (;;) {
        uint64_t delta;
        if (!gob_read_uint(p, end, pos, delta)) {
            ok = false;
            break;
        }
        if (delta == 0) break;
        field += static_cast<int>(delta);
        uint64_t u;
        if (!gob_read_uint(p, end, pos, u)) {
            ok = false;
            break;
        }
        const int64_t v = (u & 1) ? static_cast<int64_t>(~(u >> 1)) : static_cast<int64_t>(u >> 1);
        if (field >= 0 && field < args_per_call) a[field] = v;
    }
    for (int k = 0; k < args_per_call; ++k) args[j * args_per_call + k] = a[k];
    if (status) status[j] = ok ? 0 : 1;
}

}  // namespace

cudaError_t launch_resolve_names(const unsigned char* names, uint32_t stride, size_t n, ptype_method_ref* out,
                                 ptype_req* patch, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    const unsigned blocks = static_cast<unsigned>((n + kCodecThreads - 1) / kCodecThreads);
    resolve_names_kernel<<<blocks, kCodecThreads, 0, stream>>>(names, stride, n, out, patch);
    return cudaGetLastError();
}

size_t gob_scratch_bytes(size_t n) {
    const size_t blocks = (n + kCodecThreads - 1) / kCodecThreads;
    return n * sizeof(uint32_t) + (blocks + 2) * sizeof(unsigned long long);
}

cudaError_t launch_gob_encode_replies(const ptype_reply* replies, size_t n, unsigned long long* offsets,
                                      uint8_t* out, unsigned long long out_cap, void* scratch,
                                      unsigned long long* total, cudaStream_t stream) {
    if (n == 0) return cudaMemsetAsync(offsets, 0, sizeof(unsigned long long), stream);
    const size_t blocks = (n + kCodecThreads - 1) / kCodecThreads;
    unsigned long long* block_sums = static_cast<unsigned long long*>(scratch);
    uint32_t* lens = reinterpret_cast<uint32_t*>(block_sums + blocks + 2);
    const uint4* r = reinterpret_cast<const uint4*>(replies);
    gob_reply_lengths_kernel<<<static_cast<unsigned>(blocks), kCodecThreads, 0, stream>>>(r, n, lens, block_sums);
    scan_block_sums_kernel<<<1, 1024, 0, stream>>>(block_sums, blocks, total);
    gob_reply_write_kernel<<<static_cast<unsigned>(blocks), kCodecThreads, 0, stream>>>(r, n, lens, block_sums,
                                                                                          offsets, out, out_cap);
    return cudaGetLastError();
}

cudaError_t launch_gob_decode_args(const uint8_t* bytes, const unsigned long long* offsets, size_t n,
                                   int64_t* args, int args_per_call, uint8_t* status, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    const unsigned blocks = static_cast<unsigned>((n + kCodecThreads - 1) / kCodecThreads);
    gob_args_decode_kernel<<<blocks, kCodecThreads, 0, stream>>>(bytes, offsets, n, args, args_per_call, status);
    return cudaGetLastError();
}

}  // namespace ptype
