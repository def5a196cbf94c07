// dispatch.cu — the fused batched-dispatch kernels for sm_100a and their launchers.
//
// One thread handles one call end to end; one warp handles 32 consecutive calls:
//   load     LDG kernels: three coalesced 128-bit loads per lane cover the warp's 32 x 48 B
//            request records, then a 12-shuffle transpose hands every lane its own record.
//            Bulk kernels: one bulk-async (TMA) copy per tile into a per-warp shared-memory
//            ring, then three conflict-free LDS.128 per lane.
//   lookup   exact-key lookup in the 3-way cuckoo registry table, staged in shared memory by
//            bulk-async copies once per CTA (or probed in L2 when it does not fit); a per-lane
//            cache of the last service skips it for single-service batches
//            — replaces the result of Registry.Services()/nodes()
//              (cluster/registry.go:93-117,152-166)
//   route    selectNodes + hashConnectionIndex + roundRobinSelect in closed form
//            (cluster/rpc.go:176-183,252-270): with n registered nodes and
//            m = MaxConnections,  k = (m == 0 || n <= m) ? n : m,
//            r = int(seq) % k,  node = base + (k == n ? r : connHash[r] % n)
//   resolve  "Receiver.Method" the way the net/rpc server does [Go stdlib]: one perfect-hash probe
//   execute  Multiply / Add / Check (example/calculator/calculator.go:9-12,
//            example/optimus/prime.go:15-25)
//   pack     one 128-bit store per lane: 32 x 16 B contiguous reply records
//
// Kernels: dispatch_kernel (one batch, LDG), dispatch_multi_kernel and
// dispatch_multi_prefetch_kernel (up to 96 batches per launch, LDG, two tiles per warp in flight
// or one tile software-pipelined - what ptype_b200_dispatch_batches runs), dispatch_bulk_kernel /
// dispatch_bulk_multi_kernel (TMA-staged requests), call_batch_kernel (one Client's batch in
// compact form), checksum kernels.  All persistent; all carry programmatic-dependent-launch
// hooks so consecutive launches overlap ramp-up and drain.
//
// Shared memory and L1 are one 256 KiB array per SM and the streamed request loads wait in L1:
// the staged table is sized and the CTA shape chosen (plan_dispatch) so that 60 KiB of L1 remain.
//
// HBM-bound integer work: 64 algorithmic bytes per call, no tensor cores.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/ptype_b200.h"
#include "device_image.h"
#include "table_hash.h"

namespace ptype {
namespace {

constexpr int kMaxThreads = 1024;
constexpr int kSmemConn = 256;  // connection hashes kept in shared memory
constexpr uint32_t kFull = 0xFFFFFFFFu;

// ------------------------------------------------------------- PTX helpers

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}

__device__ __forceinline__ void st_stream(uint4* p, const uint4& v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// Orders this thread's earlier generic-proxy shared-memory accesses before later
// async-proxy (bulk copy) accesses to the same locations.
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// Bulk asynchronous copy global -> shared (the TMA engine, 1-D form; SASS UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ------------------------------------------------------------ arithmetic

__device__ __forceinline__ uint64_t abs_u64(int64_t v) {
    return v < 0 ? 0ull - static_cast<uint64_t>(v) : static_cast<uint64_t>(v);
}

// Prime.Check (example/optimus/prime.go:15-25), without the sleep:
//   for i := Min; i < min(Max, Target); i++ { if i != 0 && Target%i == 0 { return i } }
//   return Target
// Go's truncated % is zero exactly when |i| divides |Target| (and MinInt64 % -1 is 0,
// no trap), so the test runs on magnitudes; a 32-bit divide serves the common case.
__device__ __forceinline__ int64_t prime_check(int64_t lo, int64_t hi, int64_t target,
                                               uint64_t limit, uint32_t& status) {
    const int64_t bound = hi < target ? hi : target;
    if (lo >= bound) return target;
    const uint64_t width = static_cast<uint64_t>(bound) - static_cast<uint64_t>(lo);
    if (width > limit) {
        status = PTYPE_ST_RANGE_LIMIT;
        return 0;
    }
    if (target > 0 && target < (1ll << 22) && lo >= 1) {
        // The coordinator's streams (coordinator.go:67-73): small positive targets and
        // divisors.  One division per call, then the remainder is carried from divisor i to
        // i+1:  t = q*i + r  =>  t = q*(i+1) + (r - q),  and while r - q < 0 the quotient drops
        // by one and the remainder gains i+1.  Past sqrt(t) the quotient is below the divisor,
        // so one correction always suffices; below it the division is redone.  The division
        // itself goes through the approximate reciprocal: for t < 2^22 its quotient is off by
        // less than 1 (relative error < 2^-22) and is corrected in integers, so every
        // remainder here is exact.
        const int t = static_cast<int>(target);
        const float tf = static_cast<float>(t);
        const int end = static_cast<int>(bound);
        int i = static_cast<int>(lo);
        auto divide = [&](int d, int& q, int& r) {
            q = __float2int_rz(__fdividef(tf, static_cast<float>(d)));
            r = t - q * d;
            if (r < 0) {
                r += d;
                --q;
            } else if (r >= d) {
                r -= d;
                ++q;
            }
        };
        int q, r;
        divide(i, q, r);
        for (;;) {
            if (r == 0) return i;
            if (++i >= end) return target;
            r -= q;
            if (r < 0) {
                r += i;
                --q;
                if (r < 0) divide(i, q, r);
            }
        }
    }
    const uint64_t tmag = abs_u64(target);
    const bool t32 = (tmag >> 32) == 0;
    const uint32_t t_lo = static_cast<uint32_t>(tmag);
    for (int64_t i = lo; i < bound; ++i) {
        if (i == 0) continue;
        const uint64_t imag = abs_u64(i);
        bool divides;
        if (t32) {
            // |i| > |target| > 0 cannot divide; |target| == 0 is divisible by anything
            divides = (imag >> 32) == 0 ? (t_lo % static_cast<uint32_t>(imag)) == 0 : t_lo == 0;
        } else {
            divides = (tmag % imag) == 0;
        }
        if (divides) return i;
    }
    return target;
}

// Shared-memory state every dispatch CTA keeps next to its staged slot table.
struct CtaShared {
    ptype_method mtable[PTYPE_MTABLE_SLOTS];  // perfect-hash method table
    ptype_method methods[PTYPE_MAX_METHODS];  // linear list, error paths only
    uint32_t conn[kSmemConn];
    unsigned int status[PTYPE_ST__COUNT];
    uint64_t table_bar;
};

// The last service this lane looked up.  A batch issued by one Client names one
// service (cluster/cluster.go:101-103), so after the first tile the probe is skipped.
struct LookupCache {
    uint32_t key_lo = 0, key_hi = 0;
    uint32_t node_base = PTYPE_SLOT_EMPTY, n_and_recv = 0;
    bool valid = false;
};

// Unsigned remainder by a run-time divisor.  Kept out of line so that the common
// power-of-two case costs one AND and a not-taken branch instead of the ~25 predicated
// instructions of the inlined divide.
__device__ __noinline__ uint32_t umod_slow(uint32_t a, uint32_t d) { return a % d; }
__device__ __forceinline__ uint32_t umod(uint32_t a, uint32_t d) {
    if ((d & (d - 1u)) == 0u) return a & (d - 1u);
    return umod_slow(a, d);
}

// Round-robin index for sequence numbers that do not fit 32 bits, including the range
// where Go's int(index) is negative (cluster/rpc.go:182).
// Out-of-line helpers return (status << 32) | result so that status never needs a stack slot.
__device__ __forceinline__ uint64_t pack_result(uint32_t status, uint32_t result) {
    return (static_cast<uint64_t>(status) << 32) | result;
}

__device__ __noinline__ uint64_t route_wide(uint64_t seq, uint32_t k) {
    if (static_cast<int64_t>(seq) >= 0) return pack_result(PTYPE_ST_OK, static_cast<uint32_t>(seq % k));
    // Go's % keeps the dividend's sign: any non-zero remainder indexes clients[] below 0
    const bool bad = static_cast<int64_t>(seq) % static_cast<int64_t>(k) != 0;
    return pack_result(bad ? PTYPE_ST_BAD_INDEX : PTYPE_ST_OK, 0);
}

// Everything but the fast resolve: which net/rpc error a request gets, or the opcode of
// a method the perfect-hash entry did not confirm (cannot happen for registered methods).
__device__ __noinline__ uint64_t resolve_slow(const CtaShared& sh, uint32_t n_methods,
                                              uint32_t method, uint32_t recv, uint32_t recv_mask) {
    if (recv == PTYPE_RECV_ILL_FORMED) return pack_result(PTYPE_ST_ILL_FORMED, PTYPE_OP_NONE);
    bool recv_known = false;
    for (uint32_t e = 0; e < n_methods; ++e) {
        const ptype_method md = sh.methods[e];
        if (md.recv_hash == recv && ((recv_mask >> md.recv_bit) & 1u)) {
            recv_known = true;
            if (md.method_hash == method) return pack_result(PTYPE_ST_OK, md.opcode);
        }
    }
    return pack_result(recv_known ? PTYPE_ST_CANT_FIND_METHOD : PTYPE_ST_CANT_FIND_SERVICE,
                       PTYPE_OP_NONE);
}

// Exact-key lookup: the key sits in one of three candidate slots or nowhere.
template <bool kSmemTable>
__device__ __forceinline__ void lookup(const DeviceImage& img, const ptype_slot* slots, uint32_t key_lo,
                                       uint32_t key_hi, uint32_t& node_base, uint32_t& n_and_recv) {
    node_base = PTYPE_SLOT_EMPTY;
    n_and_recv = 0;
    if (kSmemTable) {
        // shared memory: issue all three probes at once, no divergence
        uint4 s[PTYPE_SLOT_WAYS];
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w)
            s[w] = *reinterpret_cast<const uint4*>(slots + slot_of(key_lo, key_hi, img.table_seed, w, img.n_slots));
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
            if (s[w].x == key_lo && s[w].y == key_hi && s[w].z != PTYPE_SLOT_EMPTY) {
                node_base = s[w].z;
                n_and_recv = s[w].w;
            }
        }
    } else {
        // L2: stop at the first hit so that most calls cost one sector
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
            const uint4 s = __ldg(reinterpret_cast<const uint4*>(
                slots + slot_of(key_lo, key_hi, img.table_seed, w, img.n_slots)));
            if (s.x == key_lo && s.y == key_hi && s.z != PTYPE_SLOT_EMPTY) {
                node_base = s.z;
                n_and_recv = s.w;
                break;
            }
        }
    }
}

// One call: lookup -> route -> resolve -> execute.  q0 = {svc_key, seq},
// q1 = {method, recv, a0}, q2 = {a1, a2}.
// kCache = false: look the service up every time (registries with thousands of services, where
// consecutive calls of a lane hardly ever name the same one and the compare is pure overhead).
template <bool kSmemTable, bool kCache = true>
__device__ __forceinline__ uint4 dispatch_one(const DeviceImage& img, const CtaShared& sh,
                                              const ptype_slot* slots, LookupCache& cache, uint4 q0,
                                              uint4 q1, uint4 q2) {
    if (!kCache) {
        lookup<kSmemTable>(img, slots, q0.x, q0.y, cache.node_base, cache.n_and_recv);
    } else if (!(cache.valid && cache.key_lo == q0.x && cache.key_hi == q0.y)) {
        lookup<kSmemTable>(img, slots, q0.x, q0.y, cache.node_base, cache.n_and_recv);
        cache.key_lo = q0.x;
        cache.key_hi = q0.y;
        cache.valid = true;
    }
    const uint32_t node_base = cache.node_base, n_and_recv = cache.n_and_recv;
    const uint32_t n = n_and_recv & 0xFFFFFFu;

    uint32_t status = PTYPE_ST_OK;
    uint32_t node = PTYPE_NO_NODE;
    uint64_t value = 0;
    if (node_base == PTYPE_SLOT_EMPTY) {
        status = PTYPE_ST_NO_SERVICE;  // watch delivered no nodes -> ErrNoClientAvailable
    } else if (n == 0) {
        status = PTYPE_ST_NO_CLIENT;  // rpc.go:178-180
    } else {
        // ---- route (rpc.go:176-183,252-270)
        const uint32_t m = img.max_conn;
        const bool identity = m == 0 || n <= m;  // rpc.go:253-255
        uint32_t r;
        if (q0.w == 0) {  // seq < 2^32
            // k == m is one constant per image: multiply-shift with the host's magic number
            // (Lemire fastmod); k == n differs per service and takes the divide.
            r = identity ? umod(q0.z, n) : static_cast<uint32_t>(__umul64hi(img.m_magic * q0.z, m));
        } else {
            const uint64_t rw = route_wide(
                static_cast<uint64_t>(q0.z) | (static_cast<uint64_t>(q0.w) << 32), identity ? n : m);
            r = static_cast<uint32_t>(rw);
            status = static_cast<uint32_t>(rw >> 32);
        }
        if (status == PTYPE_ST_OK) {
            uint32_t local = r;
            if (!identity) {
                const uint32_t h = img.max_conn <= kSmemConn ? sh.conn[r] : __ldg(img.conn_hash + r);
                local = umod(h, n);  // rpc.go:258-261,266-270
            }
            node = node_base + local;
            // ---- resolve [net/rpc server.readRequestHeader]: one perfect-hash probe
            const uint32_t recv_mask = n_and_recv >> 24;
            const uint4 md = *reinterpret_cast<const uint4*>(&sh.mtable[mtable_index(q1.x, img.method_mul)]);
            uint32_t opcode = md.z;
            if (!(md.x == q1.x && md.y == q1.y && opcode != PTYPE_OP_NONE && ((recv_mask >> md.w) & 1u))) {
                const uint64_t rs = resolve_slow(sh, img.n_methods, q1.x, q1.y, recv_mask);
                opcode = static_cast<uint32_t>(rs);
                status = static_cast<uint32_t>(rs >> 32);
            }
            // ---- execute
            if (status == PTYPE_ST_OK) {
                const uint64_t a0 = static_cast<uint64_t>(q1.z) | (static_cast<uint64_t>(q1.w) << 32);
                const uint64_t a1 = static_cast<uint64_t>(q2.x) | (static_cast<uint64_t>(q2.y) << 32);
                if (opcode == PTYPE_OP_CHECK) {
                    const uint64_t a2 =
                        static_cast<uint64_t>(q2.z) | (static_cast<uint64_t>(q2.w) << 32);
                    value = static_cast<uint64_t>(prime_check(static_cast<int64_t>(a0),
                                                              static_cast<int64_t>(a1),
                                                              static_cast<int64_t>(a2), img.check_limit,
                                                              status));
                } else {
                    // Go int wraps (calculator.go:10); both results computed, one selected:
                    // a mixed Add/Multiply warp does not diverge
                    const uint64_t prod = a0 * a1, sum = a0 + a1;
                    value = opcode == PTYPE_OP_MULTIPLY ? prod : sum;
                }
            }
        }
    }
    uint4 out;
    out.x = static_cast<uint32_t>(value);
    out.y = static_cast<uint32_t>(value >> 32);
    out.z = node;
    out.w = status;
    return out;
}

// Start the bulk copy of the slot table, then stage method tables / connection hashes with plain
// loads.  Ends with the table copy in flight (wait on sh.table_bar, parity 0).
template <bool kSmemTable>
__device__ __forceinline__ const ptype_slot* stage_tables(const DeviceImage& img, CtaShared& sh,
                                                          unsigned char* table_smem) {
    // The slot table's bulk copy goes out first: only thread 0 touches the barrier before the
    // __syncthreads below, so nothing has to wait for the small tables' loads (one memory latency
    // less at the head of every CTA than issuing the copy after them).
    if (kSmemTable && threadIdx.x == 0) {
        mbar_init(&sh.table_bar, 1);
        const uint32_t bytes = img.n_slots * static_cast<uint32_t>(sizeof(ptype_slot));
        mbar_expect_tx(&sh.table_bar, bytes);
        constexpr uint32_t kChunk = 32768;
        for (uint32_t off = 0; off < bytes; off += kChunk) {
            const uint32_t len = bytes - off < kChunk ? bytes - off : kChunk;
            bulk_g2s(table_smem + off, reinterpret_cast<const unsigned char*>(img.slots) + off, len,
                     &sh.table_bar);
        }
    }
    for (uint32_t i = threadIdx.x; i < PTYPE_MTABLE_SLOTS; i += blockDim.x) sh.mtable[i] = img.mtable[i];
    for (uint32_t i = threadIdx.x; i < img.n_methods; i += blockDim.x) sh.methods[i] = img.methods[i];
    if (img.max_conn <= kSmemConn)
        for (uint32_t i = threadIdx.x; i < img.max_conn; i += blockDim.x) sh.conn[i] = img.conn_hash[i];
    if (threadIdx.x < PTYPE_ST__COUNT) sh.status[threadIdx.x] = 0;
    __syncthreads();
    return kSmemTable ? reinterpret_cast<const ptype_slot*>(table_smem) : img.slots;
}

// dispatch_ex outputs for one tile: ballot/popc compaction of the failed calls (one
// global atomic per warp) and per-status counters.
__device__ __forceinline__ void tile_extras(CtaShared& sh, uint32_t lane, bool live, uint32_t status,
                                            size_t idx, uint32_t idx_base, uint32_t* failed_idx,
                                            uint32_t failed_cap, unsigned long long* stats,
                                            unsigned int& ok_calls) {
    const bool failed = live && status != PTYPE_ST_OK;
    const uint32_t live_mask = __ballot_sync(kFull, live);
    const uint32_t fail_mask = __ballot_sync(kFull, failed);
    if (lane == 0) ok_calls += __popc(live_mask & ~fail_mask);
    if (fail_mask) {
        unsigned long long base = 0;
        const int leader = __ffs(fail_mask) - 1;
        if (static_cast<int>(lane) == leader)
            base = atomicAdd(stats + kStatFailed, static_cast<unsigned long long>(__popc(fail_mask)));
        base = __shfl_sync(kFull, base, leader);
        if (failed) {
            atomicAdd(&sh.status[status], 1u);
            const unsigned long long slot = base + __popc(fail_mask & ((1u << lane) - 1u));
            if (failed_idx != nullptr && slot < failed_cap)
                failed_idx[slot] = idx_base + static_cast<uint32_t>(idx);
        }
    }
}

__device__ __forceinline__ void flush_extras(CtaShared& sh, uint32_t lane, unsigned int ok_calls,
                                             unsigned long long* stats) {
    if (lane == 0 && ok_calls) atomicAdd(&sh.status[PTYPE_ST_OK], ok_calls);
    __syncthreads();
    if (threadIdx.x < PTYPE_ST__COUNT && sh.status[threadIdx.x]) {
        atomicAdd(stats + kStatByStatus + threadIdx.x,
                  static_cast<unsigned long long>(sh.status[threadIdx.x]));
        atomicAdd(stats + kStatCalls, static_cast<unsigned long long>(sh.status[threadIdx.x]));
    }
}

// Programmatic dependent launch: let the next dispatch on the stream start filling SMs
// while this one drains (batches are independent), and do not finish before the
// previous one has, so stream order still means completion order.
__device__ __forceinline__ void griddep_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- the LDG kernels' transpose.  Lane L loads 16-byte chunks L, 32+L, 64+L of the warp's 96
// (three fully coalesced 512-byte requests); record r owns chunks 3r..3r+2.  In round c, lane s
// feeds the lane whose record needs chunk index == s (mod 32): r = 11*(s-c) mod 32
// (11 = 3^-1 mod 32), out of its register (3r+c)/32.
struct TransposePlan {
    uint32_t send_reg[3];
};
__device__ __forceinline__ TransposePlan transpose_plan(uint32_t lane) {
    TransposePlan p;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const uint32_t r = ((lane - c) * 11u) & 31u;
        p.send_reg[c] = (3u * r + c) >> 5;
    }
    return p;
}
// ld[0..2]: what this lane loaded; q[0..2]: the 48-byte record of call (tile, lane).  12 shuffles.
__device__ __forceinline__ void transpose_tile(const TransposePlan& p, uint32_t lane, const uint4 (&ld)[3],
                                               uint4 (&q)[3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const uint32_t sr = p.send_reg[c];
        const uint4 snd = sr == 0 ? ld[0] : (sr == 1 ? ld[1] : ld[2]);
        const uint32_t src = (3u * lane + c) & 31u;
        q[c].x = __shfl_sync(kFull, snd.x, src);
        q[c].y = __shfl_sync(kFull, snd.y, src);
        q[c].z = __shfl_sync(kFull, snd.z, src);
        q[c].w = __shfl_sync(kFull, snd.w, src);
    }
}

// The same data movement with the sender's choice written as selects.  Which register a lane
// sends depends on the lane modulo 3 only: chunk 32k+s belongs to round (2k+s) mod 3, so round c
// sends register k = 2(c-s) mod 3, which is 0 when s = c and 1 when s = c+1 (mod 3).  45
// instructions where the form above compiles to divergent branches around whole-uint4 moves (64),
// and one register of state instead of three.  The pipelined loop uses it; in the unrolled loops
// it measured slower (mesh image, 2000 batches: 9.97 vs 9.06 us per batch), so they keep the other.
__device__ __forceinline__ uint32_t pick3(bool first, bool second, uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t bc = second ? b : c;
    return first ? a : bc;
}
__device__ __forceinline__ void transpose_tile_sel(uint32_t lane_mod3, uint32_t lane, const uint4 (&ld)[3],
                                                   uint4 (&q)[3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const bool r0 = lane_mod3 == static_cast<uint32_t>(c), r1 = lane_mod3 == static_cast<uint32_t>((c + 1) % 3);
        const uint32_t src = (3u * lane + c) & 31u;
        q[c].x = __shfl_sync(kFull, pick3(r0, r1, ld[0].x, ld[1].x, ld[2].x), src);
        q[c].y = __shfl_sync(kFull, pick3(r0, r1, ld[0].y, ld[1].y, ld[2].y), src);
        q[c].z = __shfl_sync(kFull, pick3(r0, r1, ld[0].z, ld[1].z, ld[2].z), src);
        q[c].w = __shfl_sync(kFull, pick3(r0, r1, ld[0].w, ld[1].w, ld[2].w), src);
    }
}

// ------------------------------------------------------- kernel, bulk-staged
//
// Requests reach shared memory through the bulk-async copy engine: each warp owns a
// ring of kStages 1536-byte stages (32 records); lane 0 arms the stage's mbarrier and
// issues one cp.async.bulk per tile, every lane then reads its own 48-byte record with
// three conflict-free LDS.128 (lanes 0-7 of a quarter-warp land on 8 distinct 16-byte
// bank groups because 48/16 = 3 is odd).  No load instruction per call is spent on
// global memory and nothing is transposed through registers.
constexpr int kTileBytes = 32 * 48;
constexpr int kBulkThreads = 256;

template <bool kSmemTable, bool kExtras, int kStages>
__global__ void __launch_bounds__(kBulkThreads, 5)
    dispatch_bulk_kernel(const DeviceImage img, const uint4* __restrict__ reqs, const size_t n,
                         uint4* __restrict__ replies, uint32_t* __restrict__ failed_idx,
                         const uint32_t failed_cap, const uint32_t idx_base,
                         unsigned long long* __restrict__ stats, const uint32_t table_smem_bytes) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(16) CtaShared sh;
    __shared__ __align__(8) uint64_t s_full[(kBulkThreads / 32) * kStages];

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    unsigned char* const my_stage = dyn_smem + table_smem_bytes + warp * (kStages * kTileBytes);
    uint64_t* const my_bar = s_full + warp * kStages;

    const size_t n_tiles = (n + 31) >> 5;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * warps_per_cta;
    const size_t warp_global = static_cast<size_t>(blockIdx.x) * warps_per_cta + warp;
    const uint32_t my_tiles =
        warp_global < n_tiles
            ? static_cast<uint32_t>((n_tiles - warp_global + total_warps - 1) / total_warps)
            : 0u;

    auto issue = [&](uint32_t i, uint32_t s) {  // lane 0 only
        const size_t first = (warp_global + static_cast<size_t>(i) * total_warps) << 5;
        const size_t left = n - first;
        const uint32_t bytes = static_cast<uint32_t>(left < 32 ? left : 32) * 48u;
        mbar_expect_tx(&my_bar[s], bytes);
        bulk_g2s(my_stage + s * kTileBytes, reqs + first * 3, bytes, &my_bar[s]);
    };

    // ---- request ring first (its latency is the longest), then the tables
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(&my_bar[s], 1);
#pragma unroll
        for (int s = 0; s < kStages; ++s)
            if (static_cast<uint32_t>(s) < my_tiles) issue(s, s);
    }
    const ptype_slot* slots = stage_tables<kSmemTable>(img, sh, dyn_smem);  // has a __syncthreads
    griddep_launch_dependents();

    LookupCache cache;
    unsigned int ok_calls = 0;
    if (kSmemTable && my_tiles) mbar_wait(&sh.table_bar, 0);

    size_t idx = (warp_global << 5) + lane;          // this lane's call in tile i
    const size_t idx_step = total_warps << 5;
    for (uint32_t i = 0; i < my_tiles; ++i, idx += idx_step) {
        const uint32_t s = i % kStages;
        mbar_wait(&my_bar[s], (i / kStages) & 1u);
        const uint4* rec = reinterpret_cast<const uint4*>(my_stage + s * kTileBytes + lane * 48u);
        const uint4 q0 = rec[0], q1 = rec[1], q2 = rec[2];
        // The refill below is written by the async proxy; these reads went through the
        // generic proxy.  Without the proxy fence the two are unordered even across
        // __syncwarp, and a refill can land under a lane that has not read yet (seen on
        // B200 as records from tile i+2 answering tile i).
        fence_proxy_async();
        __syncwarp();  // every lane has its record: the stage may be refilled
        if (lane == 0 && i + kStages < my_tiles) issue(i + kStages, s);
        const bool live = idx < n;
        uint4 out = make_uint4(0, 0, 0, 0);
        if (live) {
            out = dispatch_one<kSmemTable>(img, sh, slots, cache, q0, q1, q2);
            st_stream(replies + idx, out);
        }
        if (kExtras)
            tile_extras(sh, lane, live, out.w, idx, idx_base, failed_idx, failed_cap, stats, ok_calls);
    }
    if (kSmemTable && !my_tiles) mbar_wait(&sh.table_bar, 0);  // leave no bulk copy in flight
    if (kExtras) flush_extras(sh, lane, ok_calls, stats);
    griddep_wait();
}

// ------------------------------------------------------ kernel, LDG-staged
//
// Used when the slot table needs (nearly) all of shared memory, so no room is left for
// request stages: each lane issues three coalesced 128-bit loads covering the warp's
// contiguous 1536 bytes, then a 12-shuffle transpose hands every lane its own record.
template <bool kSmemTable, bool kExtras, int kUnroll, int kThreads, int kMinCtas>
__global__ void __launch_bounds__(kThreads, kMinCtas)
    dispatch_kernel(const DeviceImage img, const uint4* __restrict__ reqs, const size_t n,
                    uint4* __restrict__ replies, uint32_t* __restrict__ failed_idx,
                    const uint32_t failed_cap, const uint32_t idx_base,
                    unsigned long long* __restrict__ stats) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(16) CtaShared sh;

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_cta = blockDim.x >> 5;

    const ptype_slot* slots = stage_tables<kSmemTable>(img, sh, dyn_smem);
    griddep_launch_dependents();

    const TransposePlan plan = transpose_plan(lane);

    const size_t n_tiles = (n + 31) >> 5;
    const size_t n_chunks = n * 3;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * warps_per_cta;
    const size_t warp_global = static_cast<size_t>(blockIdx.x) * warps_per_cta + warp;

    LookupCache cache;
    unsigned int ok_calls = 0;  // lane 0 only (kExtras)
    bool table_ready = !kSmemTable;

    for (size_t t0 = warp_global; t0 < n_tiles; t0 += total_warps * kUnroll) {
        uint4 ld[kUnroll][3];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t t = t0 + static_cast<size_t>(u) * total_warps;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const size_t c = t * 96 + j * 32 + lane;
                ld[u][j] = c < n_chunks ? ld_stream(reqs + c) : make_uint4(0, 0, 0, 0);
            }
        }
        if (!table_ready) {  // first tile's loads are already in flight
            mbar_wait(&sh.table_bar, 0);
            table_ready = true;
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t t = t0 + static_cast<size_t>(u) * total_warps;
            if (t >= n_tiles) break;  // warp-uniform
            uint4 q[3];
            transpose_tile(plan, lane, ld[u], q);
            const size_t idx = (t << 5) + lane;
            const bool live = idx < n;
            uint4 out = make_uint4(0, 0, 0, 0);
            if (live) {
                out = dispatch_one<kSmemTable>(img, sh, slots, cache, q[0], q[1], q[2]);
                st_stream(replies + idx, out);
            }
            if (kExtras)
                tile_extras(sh, lane, live, out.w, idx, idx_base, failed_idx, failed_cap, stats, ok_calls);
        }
    }
    if (!table_ready) mbar_wait(&sh.table_bar, 0);  // never leave a bulk copy in flight behind us
    if (kExtras) flush_extras(sh, lane, ok_calls, stats);
    griddep_wait();
}

// ------------------------------------------- kernel, many batches per launch
//
// ptype_b200_dispatch_batches: the tiles of up to kMaxBatchesPerLaunch batches form one
// tile stream that the persistent grid walks without ever draining - no launch gap, no
// ramp between batches, and one launch per ~hundred batches on the host side.
constexpr int kMaxBatchesPerLaunch = 96;
struct BatchList {
    const uint4* reqs[kMaxBatchesPerLaunch];
    uint4* replies[kMaxBatchesPerLaunch];
    uint32_t n[kMaxBatchesPerLaunch];
    uint32_t tile_end[kMaxBatchesPerLaunch];  // running total of 32-call tiles (< 2^32 per launch)
    int count;
    // Sequence numbers kept beside the records (ptype_b200_assign_seq_array): when set, call j of
    // the launch's ONE batch takes seq[j] instead of its record's seq field.
    const unsigned long long* seq;
    // A tail launch (launch_dispatch_multi): its CTAs are short and must not sit on an SM waiting
    // for the main launch to end, so only CTA 0 does; that is enough to keep completion in
    // stream order, because a grid is complete when all of its CTAs are.
    int tail;
};

template <bool kSmemTable, int kUnroll, int kThreads, int kMinCtas>
__global__ void __launch_bounds__(kThreads, kMinCtas)
    dispatch_multi_kernel(const DeviceImage img, const __grid_constant__ BatchList list) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(16) CtaShared sh;

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const ptype_slot* slots = stage_tables<kSmemTable>(img, sh, dyn_smem);
    griddep_launch_dependents();

    const TransposePlan plan = transpose_plan(lane);
    const uint32_t total_tiles = list.tile_end[list.count - 1];
    const uint32_t total_warps = gridDim.x * warps_per_cta;
    const uint32_t warp_global = blockIdx.x * warps_per_cta + (threadIdx.x >> 5);

    LookupCache cache;
    bool table_ready = !kSmemTable;
    int cursor[kUnroll];  // batch holding each in-flight tile; tiles only move forward
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) cursor[u] = 0;

    for (uint32_t g0 = warp_global; g0 < total_tiles; g0 += total_warps * kUnroll) {
        uint4 ld[kUnroll][3];
        uint32_t tile[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const uint32_t g = g0 + u * total_warps;
            tile[u] = 0;
            ld[u][0] = ld[u][1] = ld[u][2] = make_uint4(0, 0, 0, 0);
            if (g < total_tiles) {
                int b = cursor[u];
                while (g >= list.tile_end[b]) ++b;
                cursor[u] = b;
                tile[u] = g - (b ? list.tile_end[b - 1] : 0u);
                const size_t n_chunks = static_cast<size_t>(list.n[b]) * 3;
                const uint4* reqs = list.reqs[b];
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const size_t c = static_cast<size_t>(tile[u]) * 96 + j * 32 + lane;
                    if (c < n_chunks) ld[u][j] = ld_stream(reqs + c);
                }
            }
        }
        if (!table_ready) {
            mbar_wait(&sh.table_bar, 0);
            table_ready = true;
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            if (g0 + u * total_warps >= total_tiles) break;  // warp-uniform
            uint4 q[3];
            transpose_tile(plan, lane, ld[u], q);
            const int b = cursor[u];
            const size_t idx = (static_cast<size_t>(tile[u]) << 5) + lane;
            if (idx < list.n[b])
                st_stream(list.replies[b] + idx, dispatch_one<kSmemTable>(img, sh, slots, cache, q[0], q[1], q[2]));
        }
    }
    if (!table_ready) mbar_wait(&sh.table_bar, 0);
    if (!list.tail || blockIdx.x == 0) griddep_wait();
}

// Software-pipelined form of dispatch_multi_kernel: a warp holds ONE tile at a time.  The transpose
// empties the three load registers into the lane's own record, the next tile's loads go straight
// back into them and are in flight while this tile is dispatched.  With the 32 warps of a
// 1024-thread CTA the unrolled kernel above spent 29 % of its warp-time at the first use of a loaded
// register (profiles/r02_c3_multi_unrolled_sass_top.txt): its loads were in flight only while the
// warp waited for them.
// What a warp needs to know about its current batch (whole tiles, request and reply base with the
// lane's offset folded in) lives in registers and changes only when the warp crosses into another
// batch, ~7 tiles apart - no constant-bank table walk per tile.  The reply address is made when the
// tile is fetched; its bit 0 marks a lane beyond the batch's end.  192 instructions per tile on C3
// (the first pipelined form, which copied one register set into another: 242).
struct TileRegs {
    uint4 ld[3];
    uint4* out;              // this lane's reply record; bit 0 set: no call behind this lane
    unsigned long long seq;  // kSideSeq only
};

template <bool kSmemTable, int kThreads, int kMinCtas, bool kSideSeq>
__global__ void __launch_bounds__(kThreads, kMinCtas)
    dispatch_multi_prefetch_kernel(const DeviceImage img, const __grid_constant__ BatchList list) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(16) CtaShared sh;

    const uint32_t lane = threadIdx.x & 31u;
    constexpr uint32_t kWarps = kThreads / 32;  // launched with exactly kThreads threads

    // Where the warp is: batch b, tile `tile` of it.  b_whole = whole tiles in batch b; request and
    // reply base already point at this lane's 16 bytes of tile 0.
    int b = 0;
    uint32_t tile = blockIdx.x * kWarps + (threadIdx.x >> 5);
    uint32_t b_whole = list.n[0] >> 5;
    const uint4* b_reqs = list.reqs[0] + lane;
    uint4* b_out = list.replies[0] + lane;

    // Loads tile (b, tile) after moving (b, tile) forward into the batch that holds it; leaves
    // b == list.count when the stream has ended.
    auto fetch = [&](TileRegs& t) {
        uint32_t calls_here = 32;
        if (tile >= b_whole) {  // warp-uniform and rare: a short last tile, or on into a later batch
            for (;;) {
                const uint32_t nb = list.n[b], tiles_b = (nb + 31u) >> 5;
                if (tile < tiles_b) {
                    calls_here = nb & 31u;
                    break;
                }
                tile -= tiles_b;
                if (++b == list.count) return;
                b_whole = list.n[b] >> 5;
                b_reqs = list.reqs[b] + lane;
                b_out = list.replies[b] + lane;
                if (tile < b_whole) break;
            }
        }
        const uint4* at = b_reqs + static_cast<size_t>(tile) * 96;
        t.out = b_out + (static_cast<size_t>(tile) << 5);
        if (calls_here == 32) {
#pragma unroll
            for (int j = 0; j < 3; ++j) t.ld[j] = ld_stream(at + j * 32);
        } else {
#pragma unroll
            for (int j = 0; j < 3; ++j)
                t.ld[j] = j * 32 + lane < calls_here * 3u ? ld_stream(at + j * 32) : make_uint4(0, 0, 0, 0);
            if (lane >= calls_here) t.out = reinterpret_cast<uint4*>(reinterpret_cast<uintptr_t>(t.out) | 1u);
        }
        if (kSideSeq)  // one batch per launch in this form: call index = tile * 32 + lane
            t.seq = lane < calls_here ? __ldg(list.seq + (static_cast<size_t>(tile) << 5) + lane) : 0ull;
    };

    TileRegs t;
    fetch(t);  // before the tables: the longest latency first

    const ptype_slot* slots = stage_tables<kSmemTable>(img, sh, dyn_smem);
    griddep_launch_dependents();
    const uint32_t lane_mod3 = lane % 3u;
    LookupCache cache;
    if (kSmemTable) mbar_wait(&sh.table_bar, 0);

    while (b < list.count) {
        // The transpose empties the load registers, the next tile's loads go straight back into
        // them and are in flight while this tile is dispatched.
        uint4 q[3];
        transpose_tile_sel(lane_mod3, lane, t.ld, q);
        uint4* const out = t.out;
        if (kSideSeq) {
            q[0].z = static_cast<uint32_t>(t.seq);
            q[0].w = static_cast<uint32_t>(t.seq >> 32);
        }
        tile += gridDim.x * kWarps;
        fetch(t);
        // the 1024-thread shape means a table beyond 45 KiB: over a thousand services, no lookup cache
        if ((reinterpret_cast<uintptr_t>(out) & 1u) == 0)
            st_stream(out, dispatch_one<kSmemTable, kThreads != kMaxThreads>(img, sh, slots, cache, q[0], q[1], q[2]));
    }
    if (!list.tail || blockIdx.x == 0) griddep_wait();
}

// Bulk-staged twin of dispatch_multi_kernel: per-warp TMA ring over the tile stream of many
// batches.  Fewer registers (more resident warps) than the LDG kernel, which pays off when the
// method bodies are ALU-heavy (Prime.Check).
template <bool kSmemTable, int kStages>
__global__ void __launch_bounds__(kBulkThreads, 5)
    dispatch_bulk_multi_kernel(const DeviceImage img, const __grid_constant__ BatchList list,
                               const uint32_t table_smem_bytes) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(16) CtaShared sh;
    __shared__ __align__(8) uint64_t s_full[(kBulkThreads / 32) * kStages];

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    unsigned char* const my_stage = dyn_smem + table_smem_bytes + warp * (kStages * kTileBytes);
    uint64_t* const my_bar = s_full + warp * kStages;

    const uint32_t total_tiles = list.tile_end[list.count - 1];
    const uint32_t total_warps = gridDim.x * warps_per_cta;
    const uint32_t warp_global = blockIdx.x * warps_per_cta + warp;
    const uint32_t my_tiles = warp_global < total_tiles ? (total_tiles - warp_global + total_warps - 1) / total_warps : 0u;

    int issue_cursor = 0;  // lane 0: batch of the tile being fetched
    auto issue = [&](uint32_t i, uint32_t s) {  // lane 0 only
        const uint32_t g = warp_global + i * total_warps;
        int b = issue_cursor;
        while (g >= list.tile_end[b]) ++b;
        issue_cursor = b;
        const uint32_t tile = g - (b ? list.tile_end[b - 1] : 0u);
        const uint32_t left = list.n[b] - (tile << 5);
        const uint32_t bytes = (left < 32u ? left : 32u) * 48u;
        mbar_expect_tx(&my_bar[s], bytes);
        bulk_g2s(my_stage + s * kTileBytes, list.reqs[b] + static_cast<size_t>(tile) * 96, bytes, &my_bar[s]);
    };

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(&my_bar[s], 1);
#pragma unroll
        for (int s = 0; s < kStages; ++s)
            if (static_cast<uint32_t>(s) < my_tiles) issue(s, s);
    }
    const ptype_slot* slots = stage_tables<kSmemTable>(img, sh, dyn_smem);
    griddep_launch_dependents();

    LookupCache cache;
    if (kSmemTable && my_tiles) mbar_wait(&sh.table_bar, 0);
    int cursor = 0;
    for (uint32_t i = 0; i < my_tiles; ++i) {
        const uint32_t s = i % kStages;
        mbar_wait(&my_bar[s], (i / kStages) & 1u);
        const uint4* rec = reinterpret_cast<const uint4*>(my_stage + s * kTileBytes + lane * 48u);
        const uint4 q0 = rec[0], q1 = rec[1], q2 = rec[2];
        fence_proxy_async();  // see dispatch_bulk_kernel
        __syncwarp();
        if (lane == 0 && i + kStages < my_tiles) issue(i + kStages, s);
        const uint32_t g = warp_global + i * total_warps;
        while (g >= list.tile_end[cursor]) ++cursor;
        const uint32_t tile = g - (cursor ? list.tile_end[cursor - 1] : 0u);
        const size_t idx = (static_cast<size_t>(tile) << 5) + lane;
        if (idx < list.n[cursor])
            st_stream(list.replies[cursor] + idx, dispatch_one<kSmemTable>(img, sh, slots, cache, q0, q1, q2));
    }
    if (kSmemTable && !my_tiles) mbar_wait(&sh.table_bar, 0);
    griddep_wait();
}

// ------------------------------------------------ kernel, one Client's batch
//
// What (*Client).CallBatch submits: every call names the same service and the batch owns
// the consecutive sequence numbers seq_base+1 .. seq_base+n (one atomic add, rpc.go:181),
// so only the arguments (16 or 24 bytes) and, for mixed batches, a one-byte method
// selector travel per call.  Same lookup/route/resolve/execute as the general kernel.
struct ClientBatch {
    uint64_t svc_key;
    uint64_t seq_base;
    uint32_t index_base;  // added to the indices written to the failed list (chunked host path)
    uint32_t pad;
    uint32_t method[PTYPE_BATCH_METHODS];
    uint32_t recv[PTYPE_BATCH_METHODS];
};

template <bool kSmemTable, int kArgs, bool kHasSel, int kUnroll, int kThreads, int kMinCtas>
__global__ void __launch_bounds__(kThreads, kMinCtas)
    call_batch_kernel(const DeviceImage img, const ClientBatch batch, const int64_t* __restrict__ args,
                      const uint8_t* __restrict__ method_sel, const size_t n, uint4* __restrict__ replies,
                      unsigned long long* __restrict__ values, uint2* __restrict__ failed,
                      const uint32_t failed_cap, unsigned long long* __restrict__ failed_count,
                      unsigned long long* __restrict__ done_ctas, unsigned long long* __restrict__ host_count) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(16) CtaShared sh;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const ptype_slot* slots = stage_tables<kSmemTable>(img, sh, dyn_smem);
    griddep_launch_dependents();

    const size_t n_tiles = (n + 31) >> 5;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * warps_per_cta;
    const size_t warp_global = static_cast<size_t>(blockIdx.x) * warps_per_cta + (threadIdx.x >> 5);
    LookupCache cache;
    bool table_ready = !kSmemTable;
    const uint32_t key_lo = static_cast<uint32_t>(batch.svc_key), key_hi = static_cast<uint32_t>(batch.svc_key >> 32);

    // A warp takes groups of kGroupTiles consecutive tiles (128 calls), kUnroll tiles at a time.
    // The group's method bytes are one 128-byte line: each lane loads four of them with one
    // 32-bit load and the byte of call (tile u, lane l) is fetched from lane u*8 + l/4 by
    // shuffle - one memory request per 128 calls instead of one 32-byte request per tile,
    // which matters when the batch is streamed from host memory over PCIe.
    constexpr int kGroupTiles = 4;
    static_assert(kGroupTiles % kUnroll == 0, "a group is a whole number of unrolled rounds");
    const size_t n_groups = (n_tiles + kGroupTiles - 1) / kGroupTiles;
    const bool sel_words = kHasSel && (reinterpret_cast<uintptr_t>(method_sel) & 3u) == 0;
    for (size_t g = warp_global; g < n_groups; g += total_warps) {
        const size_t group_base = g * (kGroupTiles * 32);
        uint32_t sel_word = 0;
        if (kHasSel) {
            const size_t j = group_base + lane * 4;
            if (sel_words && j + 4 <= n) {
                sel_word = __ldg(reinterpret_cast<const uint32_t*>(method_sel + j));
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (j + k < n) sel_word |= static_cast<uint32_t>(__ldg(method_sel + j + k)) << (8 * k);
            }
        }
#pragma unroll 1
        for (int round = 0; round < kGroupTiles; round += kUnroll) {
            if (group_base + static_cast<size_t>(round) * 32 >= n) break;
            uint4 a01[kUnroll];
            uint2 a2[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const size_t idx = group_base + static_cast<size_t>(round + u) * 32 + lane;
                a01[u] = make_uint4(0, 0, 0, 0);
                a2[u] = make_uint2(0, 0);
                if (idx < n) {
                    if (kArgs == 2) {
                        a01[u] = ld_stream(reinterpret_cast<const uint4*>(args) + idx);
                    } else {
                        const uint2* p = reinterpret_cast<const uint2*>(args) + idx * 3;
                        const uint2 x = __ldg(p), y = __ldg(p + 1);
                        a01[u] = make_uint4(x.x, x.y, y.x, y.y);
                        a2[u] = __ldg(p + 2);
                    }
                }
            }
            if (!table_ready) {
                mbar_wait(&sh.table_bar, 0);
                table_ready = true;
            }
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const size_t idx = group_base + static_cast<size_t>(round + u) * 32 + lane;
                uint32_t which = 0;
                if (kHasSel) {
                    const uint32_t w = __shfl_sync(kFull, sel_word, (round + u) * 8 + (lane >> 2));
                    which = (w >> ((lane & 3u) * 8)) & 0xFFu;
                    if (which >= PTYPE_BATCH_METHODS) which = 0;
                }
                if (idx >= n) continue;
                const uint64_t seq = batch.seq_base + 1 + idx;
                uint32_t m = batch.method[0], rv = batch.recv[0];
                if (kHasSel) {
#pragma unroll
                    for (int k = 1; k < PTYPE_BATCH_METHODS; ++k) {
                        if (which == static_cast<uint32_t>(k)) {
                            m = batch.method[k];
                            rv = batch.recv[k];
                        }
                    }
                }
                const uint4 q0 = make_uint4(key_lo, key_hi, static_cast<uint32_t>(seq), static_cast<uint32_t>(seq >> 32));
                const uint4 q1 = make_uint4(m, rv, a01[u].x, a01[u].y);
                const uint4 q2 = make_uint4(a01[u].z, a01[u].w, a2[u].x, a2[u].y);
                const uint4 out = dispatch_one<kSmemTable>(img, sh, slots, cache, q0, q1, q2);
                if (values == nullptr) {
                    st_stream(replies + idx, out);
                } else {
                    // values-only replies (what CallBatch's caller reads): 8 bytes per call, plus an
                    // (index, status) entry for the rare call that failed
                    values[idx] = static_cast<unsigned long long>(out.x) | (static_cast<unsigned long long>(out.y) << 32);
                    if (out.w != PTYPE_ST_OK) {
                        const unsigned long long slot = atomicAdd(failed_count, 1ull);
                        if (slot < failed_cap) failed[slot] = make_uint2(static_cast<uint32_t>(idx) + batch.index_base, out.w);
                    }
                }
            }
        }
    }
    if (!table_ready) mbar_wait(&sh.table_bar, 0);
    griddep_wait();
    if (host_count != nullptr) {
        // The CTA that finishes last hands the failure count to the host and leaves both
        // counters at zero for the next batch (batches of one context never overlap).
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(done_ctas, 1ull) == gridDim.x - 1) {
                __threadfence();
                *host_count = atomicExch(failed_count, 0ull);
                *done_ctas = 0;
                __threadfence_system();
            }
        }
    }
}

// ---------------------------------------------------------- reply checksum

__global__ void __launch_bounds__(kChecksumThreads)
    checksum_partials_kernel(const uint4* __restrict__ replies, const size_t n,
                             const uint64_t index_base, unsigned long long* __restrict__ partials) {
    unsigned long long acc = 0;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint4 r = ld_stream(replies + i);
        const uint64_t value = static_cast<uint64_t>(r.x) | (static_cast<uint64_t>(r.y) << 32);
        acc += checksum_term(value, r.z, r.w, index_base + i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
    __shared__ unsigned long long s_part[kChecksumThreads / 32];
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = 0;
        for (int w = 0; w < kChecksumThreads / 32; ++w) b += s_part[w];
        partials[blockIdx.x] = b;
    }
}

__global__ void __launch_bounds__(kChecksumThreads)
    checksum_final_kernel(const unsigned long long* __restrict__ partials, const int n_partials,
                          unsigned long long* __restrict__ out) {
    unsigned long long acc = 0;
    for (int i = threadIdx.x; i < n_partials; i += blockDim.x) acc += partials[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
    __shared__ unsigned long long s_part[kChecksumThreads / 32];
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = 0;
        for (int w = 0; w < kChecksumThreads / 32; ++w) b += s_part[w];
        *out = b;
    }
}

// LDG kernel shapes: one 1024-thread CTA per SM when the staged table is large (2 tiles in
// flight per warp, 64 registers), several 256-thread CTAs otherwise.
#ifndef PTYPE_SMALL_UNROLL
#define PTYPE_SMALL_UNROLL 2
#endif
#ifndef PTYPE_SMALL_MINCTAS
#define PTYPE_SMALL_MINCTAS 4
#endif
constexpr int kSmallThreads = 256;
constexpr int kSmallUnroll = PTYPE_SMALL_UNROLL;
constexpr int kSmallMinCtas = PTYPE_SMALL_MINCTAS;
constexpr int kStagesDefault = 2;
constexpr size_t kBulkStageBytes = (kBulkThreads / 32) * kStagesDefault * kTileBytes;  // per CTA

bool pdl_enabled() {
    static const bool on = [] {
        const char* v = std::getenv("PTYPE_B200_PDL");
        return !(v && v[0] == '0');
    }();
    return on;
}

// One launch.  `pdl` adds programmatic stream serialization, so that this kernel may start while
// its predecessor on the stream drains; the caller passes it only when that predecessor is another
// dispatch launch of this library (whose output this kernel does not read) - see after_dispatch in
// context.cu.  A kernel that asks for more than 48 KiB of dynamic shared memory has to opt in, per
// function and per device; doing it here covers every variant that is ever launched.
// The 48 KiB a kernel gets without opting in cover its STATIC shared memory too (the method tables,
// ~3 KiB), so the opt-in starts well below 48 KiB of dynamic memory: a 46.9 KiB table (1 500
// services) was refused before this.  Asking for less than the default is harmless.
constexpr size_t kOptInAbove = 32 * 1024;
template <typename Kernel, typename... Args>
cudaError_t launch_kernel(Kernel kernel, int blocks, int threads, size_t smem, cudaStream_t stream, bool pdl,
                          Args... args) {
    if (smem > kOptInAbove) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(blocks));
    cfg.blockDim = dim3(static_cast<unsigned>(threads));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

template <bool kSmemTable, bool kExtras>
cudaError_t launch_variant(const DeviceImage& img, const LaunchPlan& plan, int blocks,
                           const ptype_req* reqs, size_t n, ptype_reply* replies, uint32_t* failed_idx,
                           uint32_t failed_cap, uint32_t idx_base, unsigned long long* stats,
                           cudaStream_t stream, bool pdl) {
    const uint4* q = reinterpret_cast<const uint4*>(reqs);
    uint4* r = reinterpret_cast<uint4*>(replies);
    if (plan.bulk)
        return launch_kernel(dispatch_bulk_kernel<kSmemTable, kExtras, kStagesDefault>, blocks,
                             plan.threads, plan.dyn_smem, stream, pdl, img, q, n, r, failed_idx, failed_cap,
                             idx_base, stats, static_cast<uint32_t>(plan.table_smem));
    if (plan.threads == kMaxThreads)
        return launch_kernel(dispatch_kernel<kSmemTable, kExtras, 2, kMaxThreads, 1>, blocks,
                             plan.threads, plan.dyn_smem, stream, pdl, img, q, n, r, failed_idx, failed_cap,
                             idx_base, stats);
    return launch_kernel(dispatch_kernel<kSmemTable, kExtras, kSmallUnroll, kSmallThreads, kSmallMinCtas>,
                         blocks, plan.threads, plan.dyn_smem, stream, pdl, img, q, n, r, failed_idx,
                         failed_cap, idx_base, stats);
}

template <typename Kernel>
cudaError_t occupancy_of(Kernel kernel, int threads, size_t dyn_smem, int* ctas_per_sm) {
    if (dyn_smem > kOptInAbove) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(dyn_smem));
        if (e != cudaSuccess) return e;
    }
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, kernel, threads, dyn_smem);
}

template <bool kSmemTable, bool kExtras>
cudaError_t occupancy_variant(const LaunchPlan& plan, int* ctas_per_sm) {
    if (plan.bulk)
        return occupancy_of(dispatch_bulk_kernel<kSmemTable, kExtras, kStagesDefault>, plan.threads,
                            plan.dyn_smem, ctas_per_sm);
    if (plan.threads == kMaxThreads)
        return occupancy_of(dispatch_kernel<kSmemTable, kExtras, 2, kMaxThreads, 1>, plan.threads,
                            plan.dyn_smem, ctas_per_sm);
    return occupancy_of(dispatch_kernel<kSmemTable, kExtras, kSmallUnroll, kSmallThreads, kSmallMinCtas>,
                        plan.threads, plan.dyn_smem, ctas_per_sm);
}

}  // namespace

// Which kernel, block shape and residency (measured on B200, profiles/README.md):
//   slot table <= 200 KiB : table staged in shared memory, requests by LDG + shuffle
//                           transpose (four 256-thread CTAs per SM while four copies of the
//                           table leave 60 KiB of L1, i.e. up to 45 KiB; one 1024-thread CTA
//                           above; the image builder sizes staged tables to 192 KiB at most)
//   larger                : table probed in L2, requests staged by bulk-async copies
//   table <= 16 KiB and every exposed receiver is Prime (ALU-heavy calls): bulk-async
//                           request staging with the table in shared memory
// PTYPE_F_GLOBAL_TABLE / PTYPE_F_LDG_LOADS / PTYPE_F_BULK_LOADS force the other variants
// (bulk staging with a shared-memory table needs the table to fit 16 KiB).
// The persistent grid is ctas_per_sm x SM count (148 on B200) in every case.
cudaError_t plan_dispatch(const DeviceImage& img, uint32_t flags, bool extras, LaunchPlan* plan) {
    const size_t table_bytes = static_cast<size_t>(img.n_slots) * sizeof(ptype_slot);
    constexpr size_t kSmemBudget = 200 * 1024;
    constexpr size_t kSmallTable = 16 * 1024;
    const bool force_global = (flags & PTYPE_F_GLOBAL_TABLE) != 0;
    const bool force_ldg = (flags & PTYPE_F_LDG_LOADS) != 0;
    const bool force_bulk = (flags & PTYPE_F_BULK_LOADS) != 0 && !force_ldg;
    // An image whose services expose only Prime.Check will be sent ALU-heavy calls: the
    // bulk-staged kernel keeps more warps resident (48 registers) and measured 98 vs 76 G
    // calls/s on the optimus replay, while losing on Add/Multiply streams (67 vs 114).
    const bool alu_heavy = img.check_only != 0 && table_bytes <= kSmallTable && !force_ldg && !force_global;
    plan->smem_table = !force_global && table_bytes <= kSmemBudget &&
                       !(force_bulk && table_bytes > kSmallTable);
    plan->bulk = force_bulk || alu_heavy || (!force_ldg && !plan->smem_table);
    plan->table_smem = plan->smem_table ? (table_bytes + 127) / 128 * 128 : 0;
    if (plan->bulk) {
        plan->threads = kBulkThreads;
        plan->dyn_smem = plan->table_smem + kBulkStageBytes;
    } else {
        // Shared memory and L1 are one 256 KiB array per SM, and the streamed request loads wait in
        // L1 until they arrive.  Once the CTAs of an SM together take more than 196 KiB of shared
        // memory, 28 KiB of L1 are left, too little for the ~48 KiB the 32 warps keep in flight:
        // 11.0 instead of 10.1 us per 2^20 calls (tools/table_band_probe.py).  So four 256-thread
        // CTAs only while four copies of the table stay inside 196 KiB (8.8 us); above that one
        // 1024-thread CTA (9.5 us for any table up to the budget, where 2 x 256 threads beside a
        // 97 KiB table took 12.0).
        constexpr size_t kSmemWithRoomForL1 = 196 * 1024, kPerCtaOverhead = sizeof(CtaShared) + 1024;
        const bool four_fit = kSmallMinCtas * (plan->table_smem + kPerCtaOverhead) <= kSmemWithRoomForL1;
        plan->threads = four_fit ? kSmallThreads : kMaxThreads;
        plan->dyn_smem = plan->table_smem;
    }
    int ctas = 0;
    cudaError_t e;
    if (plan->smem_table)
        e = extras ? occupancy_variant<true, true>(*plan, &ctas) : occupancy_variant<true, false>(*plan, &ctas);
    else
        e = extras ? occupancy_variant<false, true>(*plan, &ctas) : occupancy_variant<false, false>(*plan, &ctas);
    if (e != cudaSuccess) return e;
    if (ctas < 1) return cudaErrorLaunchOutOfResources;
    plan->ctas_per_sm = ctas;
    return cudaSuccess;
}

int grid_for(const LaunchPlan& plan, size_t n, int sm_count) {
    const size_t tiles = (n + 31) / 32;
    const size_t warps_per_cta = static_cast<size_t>(plan.threads) / 32;
    size_t want = (tiles + warps_per_cta - 1) / warps_per_cta;
    const size_t resident = static_cast<size_t>(sm_count) * plan.ctas_per_sm;
    if (want > resident) want = resident;
    return want < 1 ? 1 : static_cast<int>(want);
}

cudaError_t launch_dispatch(const DeviceImage& img, const LaunchPlan& plan, int blocks,
                            const ptype_req* reqs, size_t n, ptype_reply* replies, uint32_t* failed_idx,
                            uint32_t failed_cap, uint32_t idx_base, unsigned long long* stats,
                            cudaStream_t stream, bool pdl) {
    const bool extras = stats != nullptr;
    if (plan.smem_table) {
        return extras ? launch_variant<true, true>(img, plan, blocks, reqs, n, replies, failed_idx,
                                                   failed_cap, idx_base, stats, stream, pdl)
                      : launch_variant<true, false>(img, plan, blocks, reqs, n, replies, failed_idx,
                                                    failed_cap, idx_base, stats, stream, pdl);
    }
    return extras ? launch_variant<false, true>(img, plan, blocks, reqs, n, replies, failed_idx,
                                                failed_cap, idx_base, stats, stream, pdl)
                  : launch_variant<false, false>(img, plan, blocks, reqs, n, replies, failed_idx,
                                                 failed_cap, idx_base, stats, stream, pdl);
}

// Up to kMaxBatchesPerLaunch device-resident batches in ONE launch.
int max_batches_per_launch() { return kMaxBatchesPerLaunch; }

namespace {
// Appends records [first, first + n) of a batch to a list.
void list_add(BatchList& list, const ptype_batch& bt, uint64_t first, uint64_t n) {
    const int m = list.count;
    list.reqs[m] = reinterpret_cast<const uint4*>(bt.reqs + first);
    list.replies[m] = reinterpret_cast<uint4*>(bt.replies + first);
    list.n[m] = static_cast<uint32_t>(n);
    list.tile_end[m] = (m ? list.tile_end[m - 1] : 0u) + static_cast<uint32_t>((n + 31) / 32);
    list.count = m + 1;
}

cudaError_t launch_list(const DeviceImage& img, const LaunchPlan& plan, const BatchList& list, int blocks,
                        cudaStream_t stream, bool pdl);
}  // namespace

// A lone launch ends with a tail: with equal shares the SMs of a B200 finish 4-7 % apart in the
// pipelined loop and up to 20 % in the unrolled one (tools/tail_probe.cu, tools/spread_probe.cu),
// and whoever is done early idles until the last one is.  Handing tiles
// out through counters costs more than it saves once launches follow each other
// (profiles/r02_tile_scheduling.txt), so the hardware's own block scheduler does it, for the end of
// the launch only: a list of 2..kTailMaxBatches batches with a small table is cut in two, the first
// part goes to the persistent grid, the last kTailPercent % to a second launch of short CTAs
// (kTailTilesPerWarp tiles per warp) behind it.  Those start wherever a persistent CTA has left.
// Only for a list that arrives on an idle stream (`lone`): when lists follow each other the next
// launch fills the tail by itself, and a tail launch in between costs 4-9 % (lists of 20 and 40
// batches back to back: 9.40 -> 9.80 and 9.02 -> 9.88 us per batch, tools/list_stream_probe.py).
// Longer lists (a host with that much queued keeps launches coming) and single batches (two
// launches per batch would cost the host more than the tail) stay whole as well.
constexpr int kTailMaxBatches = 48;
constexpr uint32_t kTailTilesPerWarp = 4;
constexpr uint64_t kTailMinTiles = 16384;  // lists shorter than half a 2^20-call batch are not worth a second launch

cudaError_t launch_dispatch_multi(const DeviceImage& img, const LaunchPlan& plan, int sm_count,
                                  const ptype_batch* batches, int count, cudaStream_t stream, bool pdl,
                                  const unsigned long long* seq, int* kernels, bool lone) {
    if (kernels) *kernels = 0;
    if (seq != nullptr && (count != 1 || plan.bulk)) return cudaErrorInvalidValue;
    uint64_t tiles = 0;
    int live = 0;
    for (int b = 0; b < count; ++b) {
        if (batches[b].n == 0) continue;
        if (batches[b].n > 0xFFFFFFFFull) return cudaErrorInvalidValue;  // the list holds 32-bit counts
        tiles += (batches[b].n + 31) / 32;
        ++live;
    }
    if (live == 0) return cudaSuccess;
    if (tiles >= 0xFFFFFFFFull) return cudaErrorInvalidValue;
    static const uint64_t tail_percent = [] {
        const char* e = std::getenv("PTYPE_B200_TAIL_PERCENT");
        return e ? static_cast<uint64_t>(std::atoi(e)) : uint64_t(8);
    }();
    const bool split = lone && tail_percent > 0 && tail_percent < 100 && seq == nullptr && !plan.bulk &&
                       plan.threads == kSmallThreads && live >= 2 && live <= kTailMaxBatches && tiles >= kTailMinTiles;
    const uint64_t tail_tiles = split ? tiles * tail_percent / 100 : 0;
    const uint64_t main_tiles = tiles - tail_tiles;

    BatchList main_list = {}, tail_list = {};
    main_list.seq = seq;
    tail_list.tail = 1;
    uint64_t seen = 0;  // tiles before the current batch
    for (int b = 0; b < count; ++b) {
        const uint64_t n = batches[b].n;
        if (n == 0) continue;
        const uint64_t bt = (n + 31) / 32;
        if (seen + bt <= main_tiles) {
            list_add(main_list, batches[b], 0, n);
        } else if (seen >= main_tiles) {
            list_add(tail_list, batches[b], 0, n);
        } else {  // the cut falls inside this batch, on a tile boundary
            const uint64_t head = (main_tiles - seen) * 32;
            list_add(main_list, batches[b], 0, head);
            list_add(tail_list, batches[b], head, n - head);
        }
        seen += bt;
    }
    const size_t warps_per_cta = static_cast<size_t>(plan.threads) / 32;
    const size_t want = (main_tiles + warps_per_cta - 1) / warps_per_cta;
    const size_t resident = static_cast<size_t>(sm_count) * plan.ctas_per_sm;
    const int blocks = static_cast<int>(want < resident ? (want ? want : 1) : resident);
    cudaError_t e = launch_list(img, plan, main_list, blocks, stream, pdl);
    if (e != cudaSuccess) return e;
    if (kernels) *kernels = 1;
    if (tail_list.count == 0) return cudaSuccess;
    const uint64_t per_cta = warps_per_cta * kTailTilesPerWarp;
    e = launch_list(img, plan, tail_list, static_cast<int>((tail_tiles + per_cta - 1) / per_cta), stream, true);
    if (e == cudaSuccess && kernels) *kernels = 2;
    return e;
}

namespace {
cudaError_t launch_list(const DeviceImage& img, const LaunchPlan& plan, const BatchList& list, int blocks,
                        cudaStream_t stream, bool pdl) {
    const unsigned long long* seq = list.seq;
    const uint32_t table_smem = static_cast<uint32_t>(plan.table_smem);
    if (plan.bulk) {
        if (plan.smem_table)
            return launch_kernel(dispatch_bulk_multi_kernel<true, kStagesDefault>, blocks, plan.threads,
                                 plan.dyn_smem, stream, pdl, img, list, table_smem);
        return launch_kernel(dispatch_bulk_multi_kernel<false, kStagesDefault>, blocks, plan.threads,
                             plan.dyn_smem, stream, pdl, img, list, table_smem);
    }
    // Which loop: the software-pipelined kernel wins wherever a tile costs more instructions than a
    // load takes to arrive - the 1024-thread shape (C3: 191 -> 185 us per 2^24 calls) and the
    // 256-thread shape when nodes are picked through the connection hashes (C2: 10.9 -> 10.5 us per
    // batch in a lone launch of 20, 9.31 -> 9.23 pipelined); with the identity selection of a mesh
    // the tile is cheap and the unrolled kernel's two tiles in flight per warp do better (8.70 vs
    // 9.15 us).  PTYPE_B200_PREFETCH overrides: 0 = never, 1 = the 1024-thread shape only,
    // 2 = always (measurement switch).
    static const int forced = [] {
        const char* e = std::getenv("PTYPE_B200_PREFETCH");
        return e ? std::atoi(e) : -1;
    }();
    const int prefetch = forced >= 0 ? forced : (img.max_conn != 0 ? 2 : 1);
    if (seq != nullptr) {  // sequence numbers beside the records: the pipelined loop in either shape
        if (plan.threads == kMaxThreads)
            return plan.smem_table ? launch_kernel(dispatch_multi_prefetch_kernel<true, kMaxThreads, 1, true>, blocks,
                                                   plan.threads, plan.dyn_smem, stream, pdl, img, list)
                                   : launch_kernel(dispatch_multi_prefetch_kernel<false, kMaxThreads, 1, true>, blocks,
                                                   plan.threads, plan.dyn_smem, stream, pdl, img, list);
        return plan.smem_table
                   ? launch_kernel(dispatch_multi_prefetch_kernel<true, kSmallThreads, kSmallMinCtas, true>, blocks,
                                   plan.threads, plan.dyn_smem, stream, pdl, img, list)
                   : launch_kernel(dispatch_multi_prefetch_kernel<false, kSmallThreads, kSmallMinCtas, true>, blocks,
                                   plan.threads, plan.dyn_smem, stream, pdl, img, list);
    }
    if (plan.threads == kMaxThreads) {
        if (prefetch >= 1) {
            if (plan.smem_table)
                return launch_kernel(dispatch_multi_prefetch_kernel<true, kMaxThreads, 1, false>, blocks, plan.threads,
                                     plan.dyn_smem, stream, pdl, img, list);
            return launch_kernel(dispatch_multi_prefetch_kernel<false, kMaxThreads, 1, false>, blocks, plan.threads,
                                 plan.dyn_smem, stream, pdl, img, list);
        }
        if (plan.smem_table)
            return launch_kernel(dispatch_multi_kernel<true, 2, kMaxThreads, 1>, blocks, plan.threads,
                                 plan.dyn_smem, stream, pdl, img, list);
        return launch_kernel(dispatch_multi_kernel<false, 2, kMaxThreads, 1>, blocks, plan.threads,
                             plan.dyn_smem, stream, pdl, img, list);
    }
    if (prefetch >= 2) {
        if (plan.smem_table)
            return launch_kernel(dispatch_multi_prefetch_kernel<true, kSmallThreads, kSmallMinCtas, false>, blocks,
                                 plan.threads, plan.dyn_smem, stream, pdl, img, list);
        return launch_kernel(dispatch_multi_prefetch_kernel<false, kSmallThreads, kSmallMinCtas, false>, blocks,
                             plan.threads, plan.dyn_smem, stream, pdl, img, list);
    }
    if (plan.smem_table)
        return launch_kernel(dispatch_multi_kernel<true, kSmallUnroll, kSmallThreads, kSmallMinCtas>, blocks,
                             plan.threads, plan.dyn_smem, stream, pdl, img, list);
    return launch_kernel(dispatch_multi_kernel<false, kSmallUnroll, kSmallThreads, kSmallMinCtas>, blocks,
                         plan.threads, plan.dyn_smem, stream, pdl, img, list);
}

}  // namespace

namespace {
template <bool kSmemTable, int kArgs, bool kHasSel>
cudaError_t launch_call_batch_variant(const DeviceImage& img, const LaunchPlan& plan, int blocks,
                                      const ClientBatch& batch, const int64_t* args,
                                      const uint8_t* method_sel, size_t n, ptype_reply* replies,
                                      const CallBatchOut& o, cudaStream_t stream, bool pdl) {
    uint4* r = reinterpret_cast<uint4*>(replies);
    unsigned long long* values = reinterpret_cast<unsigned long long*>(o.values);
    uint2* failed = reinterpret_cast<uint2*>(o.failed);
    if (plan.threads == kMaxThreads)
        return launch_kernel(call_batch_kernel<kSmemTable, kArgs, kHasSel, 2, kMaxThreads, 1>, blocks,
                             plan.threads, plan.dyn_smem, stream, pdl, img, batch, args, method_sel, n, r,
                             values, failed, o.failed_cap, o.failed_count, o.done_ctas, o.host_count);
    return launch_kernel(call_batch_kernel<kSmemTable, kArgs, kHasSel, 2, kSmallThreads, kSmallMinCtas>,
                         blocks, plan.threads, plan.dyn_smem, stream, pdl, img, batch, args, method_sel, n, r,
                         values, failed, o.failed_cap, o.failed_count, o.done_ctas, o.host_count);
}
}  // namespace

// One Client's batch (see call_batch_kernel).  Uses the LDG-shaped plan of the image.
cudaError_t launch_call_batch(const DeviceImage& img, const LaunchPlan& plan, int blocks,
                              uint64_t svc_key, uint64_t seq_base, const uint32_t* methods,
                              const uint32_t* recvs, int n_methods, const int64_t* args,
                              int args_per_call, const uint8_t* method_sel, size_t n,
                              ptype_reply* replies, const CallBatchOut& out, cudaStream_t stream, bool pdl) {
    ClientBatch batch = {};
    batch.svc_key = svc_key;
    batch.seq_base = seq_base;
    batch.index_base = out.index_base;
    for (int i = 0; i < PTYPE_BATCH_METHODS; ++i) {
        batch.method[i] = methods[i < n_methods ? i : 0];
        batch.recv[i] = recvs[i < n_methods ? i : 0];
    }
    const bool sel = method_sel != nullptr;
#define PTYPE_CB(SM, AR, SEL) \
    launch_call_batch_variant<SM, AR, SEL>(img, plan, blocks, batch, args, method_sel, n, replies, out, stream, pdl)
    if (plan.smem_table) {
        if (args_per_call == 2) return sel ? PTYPE_CB(true, 2, true) : PTYPE_CB(true, 2, false);
        return sel ? PTYPE_CB(true, 3, true) : PTYPE_CB(true, 3, false);
    }
    if (args_per_call == 2) return sel ? PTYPE_CB(false, 2, true) : PTYPE_CB(false, 2, false);
    return sel ? PTYPE_CB(false, 3, true) : PTYPE_CB(false, 3, false);
#undef PTYPE_CB
}

cudaError_t launch_reply_checksum(const ptype_reply* replies, size_t n, uint64_t index_base,
                                  unsigned long long* partials, unsigned long long* out,
                                  cudaStream_t stream) {
    checksum_partials_kernel<<<kChecksumBlocks, kChecksumThreads, 0, stream>>>(
        reinterpret_cast<const uint4*>(replies), n, index_base, partials);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    checksum_final_kernel<<<1, kChecksumThreads, 0, stream>>>(partials, kChecksumBlocks, out);
    return cudaGetLastError();
}

}  // namespace ptype

// ------------------------------------------------ checksum without allocation
//
// Some hosts of this library must not trigger a device allocation while they
// dispatch: cudaMalloc and cudaFree stall every stream of the device until they
// return, and a process that has handed all of its HBM to its own pool (or runs
// under a fixed memory quota) may simply be refused.  ptype_b200_reply_checksum
// lazily cudaMallocs a partial-sum buffer per context; the entry point below is
// for those hosts.  Its partial sums live in the fixed array here instead, which
// every GPU that loads this library carries (128 KiB) until the process exits.
namespace ptype {
namespace {

constexpr int kNoAllocTile = kChecksumThreads * 16;  // replies folded into one partial sum
constexpr int kNoAllocMaxPartials = 16384;           // 16384 x 8 B = 128 KiB of scratch
constexpr size_t kNoAllocMaxReplies = static_cast<size_t>(kNoAllocMaxPartials) * kNoAllocTile;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) unsigned long long g_noalloc_partials[kNoAllocMaxPartials];

// Device addresses of the array above, one per GPU.  In host code the symbol name is
// only a handle for the runtime, so the address each GPU gave the array has to be
// asked for, once, while that GPU is the current device.
unsigned long long* g_noalloc_addr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mu;

cudaError_t noalloc_scratch(int device, unsigned long long** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (g_noalloc_addr[device] == nullptr) {
        unsigned long long* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_partials);
        if (e != cudaSuccess) return e;
        g_noalloc_addr[device] = ptr;
    }
    *out = g_noalloc_addr[device];
    return cudaSuccess;
}

}  // namespace
}  // namespace ptype

extern "C" int ptype_b200_reply_checksum_noalloc(int device, const ptype_reply* d_replies, size_t n,
                                                 uint64_t index_base, uint64_t* d_checksum,
                                                 void* cuda_stream) {
    using namespace ptype;
    char msg[256];
    if (device < 0 || device >= kNoAllocMaxDevices || d_checksum == nullptr ||
        (n != 0 && d_replies == nullptr)) {
        set_last_error("ptype_b200_reply_checksum_noalloc: bad device ordinal or null pointer");
        return PTYPE_E_INVAL;
    }
    if (n > kNoAllocMaxReplies) {
        std::snprintf(msg, sizeof msg,
                      "ptype_b200_reply_checksum_noalloc: n=%zu is beyond the %zu replies the built-in "
                      "%d KiB scratch array covers; split the range or use ptype_b200_reply_checksum",
                      n, kNoAllocMaxReplies,
                      static_cast<int>(kNoAllocMaxPartials * sizeof(unsigned long long) / 1024));
        set_last_error(msg);
        return PTYPE_E_LIMIT;
    }
    int previous = -1;
    cudaError_t e = cudaGetDevice(&previous);
    if (e == cudaSuccess && previous != device) e = cudaSetDevice(device);
    unsigned long long* partials = nullptr;
    if (e == cudaSuccess) e = noalloc_scratch(device, &partials);  // first call on this GPU: set-up
    if (e == cudaSuccess) {
        cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
        const int tiles = static_cast<int>((n + kNoAllocTile - 1) / kNoAllocTile);
        // Every partial the second pass reads is written by the first pass of this same
        // call, so nothing left behind by an earlier call needs clearing.
        if (tiles > 0) {
            checksum_partials_kernel<<<tiles, kChecksumThreads, 0, stream>>>(
                reinterpret_cast<const uint4*>(d_replies), n, index_base, partials);
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) {
            checksum_final_kernel<<<1, kChecksumThreads, 0, stream>>>(
                partials, tiles, reinterpret_cast<unsigned long long*>(d_checksum));
            e = cudaGetLastError();
        }
    }
    if (previous >= 0 && previous != device) cudaSetDevice(previous);
    if (e != cudaSuccess) {
        std::snprintf(msg, sizeof msg, "ptype_b200_reply_checksum_noalloc: %s (%s)",
                      cudaGetErrorString(e), cudaGetErrorName(e));
        set_last_error(msg);
        return PTYPE_E_CUDA;
    }
    return PTYPE_OK;
}
