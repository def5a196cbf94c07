// table_hash.h — hash functions shared by the host image builder and the kernels.
// Everything here must give the same bits on host and device.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define PTYPE_HD __host__ __device__ __forceinline__
#else
#define PTYPE_HD inline
#endif

namespace ptype {

// hash/fnv New32a [Go stdlib]; the reference's one explicit hash (cluster/rpc.go:267).
PTYPE_HD uint32_t fnv1a32(const void* data, size_t len) {
    const unsigned char* s = static_cast<const unsigned char*>(data);
    uint32_t h = 0x811c9dc5u;
    for (size_t i = 0; i < len; ++i) {
        h ^= s[i];
        h *= 0x01000193u;
    }
    return h;
}

// hash/fnv New64a with the offset basis xor-ed by `salt` (0 = plain New64a).
PTYPE_HD uint64_t fnv1a64(const void* data, size_t len, uint64_t salt) {
    const unsigned char* s = static_cast<const unsigned char*>(data);
    uint64_t h = 0xcbf29ce484222325ull ^ salt;
    for (size_t i = 0; i < len; ++i) {
        h ^= s[i];
        h *= 0x100000001b3ull;
    }
    return h;
}

// murmur3 finaliser (declared here, defined below).
PTYPE_HD uint64_t fmix64(uint64_t x);

// Checksum of a registry image body (ptype_image_header.body_checksum): four FNV-1a-style
// lanes over little-endian 64-bit words (word i goes to lane i mod 4), folded at the end.
// Word-wise and four-lane so that validating a 200 KB table costs microseconds, not the
// third of a millisecond a byte-serial FNV takes - it runs on every publish.  `len` is a
// multiple of 8 (image sections are 16-byte aligned).
inline uint64_t image_body_hash(const void* data, size_t len) {
    const unsigned char* s = static_cast<const unsigned char*>(data);
    const size_t n_words = len / 8;
    uint64_t lane[4] = {0xcbf29ce484222325ull, 0xcbf29ce484222325ull ^ 1, 0xcbf29ce484222325ull ^ 2,
                        0xcbf29ce484222325ull ^ 3};
    for (size_t i = 0; i < n_words; ++i) {
        uint64_t w;
        memcpy(&w, s + 8 * i, 8);  // hosts this library runs on are little-endian
        lane[i & 3] = (lane[i & 3] ^ w) * 0x100000001b3ull;
    }
    uint64_t h = n_words;
    for (int j = 0; j < 4; ++j) h = fmix64(h ^ lane[j]) + 0x9E3779B97F4A7C15ull;
    return h;
}

// murmur3 finaliser: spreads the FNV key before it picks a slot / mixes a checksum.
PTYPE_HD uint64_t fmix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

// Candidate slot `way` (0..2) of a key in a table of n_slots (any n_slots >= 1).  The key
// is already an FNV-1a-64 hash: xor-fold it to 32 bits, mix in the table seed, scramble
// with one odd multiply per way, then reduce by multiply-shift, so the table need not be
// a power of two and can be sized to fill shared memory.  Three integer instructions per
// way on the device.
PTYPE_HD uint32_t slot_of(uint32_t key_lo, uint32_t key_hi, uint32_t seed, int way, uint32_t n_slots) {
    const uint32_t mul = way == 0 ? 0x9E3779B1u : (way == 1 ? 0x85EBCA77u : 0xC2B2AE3Du);
    const uint32_t h = ((key_lo ^ key_hi) ^ seed) * mul;
    return static_cast<uint32_t>((static_cast<uint64_t>(h) * n_slots) >> 32);
}

// Index into the 64-entry perfect-hash method table.
PTYPE_HD uint32_t mtable_index(uint32_t method_hash, uint32_t mul) { return (method_hash * mul) >> 26; }

// One term of the order-independent reply checksum (see ptype_b200_reply_checksum).
PTYPE_HD uint64_t checksum_term(uint64_t value, uint32_t node, uint32_t status, uint64_t idx) {
    uint64_t w1 = static_cast<uint64_t>(node) | (static_cast<uint64_t>(status) << 32);
    return fmix64(value + 0x9E3779B97F4A7C15ull * (idx + 1)) ^
           fmix64(w1 ^ (idx * 0xC2B2AE3D27D4EB4Full + 0x165667B19E3779F9ull));
}

}  // namespace ptype
