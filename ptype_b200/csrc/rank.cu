// rank.cu — the balancer's counters on the device.
//
// Every Client of the reference owns one counter; each call that finds a non-empty client list
// takes its next value (cluster/rpc.go:129,172-183: `if len(clients) == 0 { return nil }` comes
// BEFORE `atomic.AddUint64(&c.seq, 1)`, so calls to a service without nodes consume nothing).
// A batch that mixes services is N such calls in batch order, so call j gets
//     seq_j = counter[service_j] + 1 + #{ i < j : service_i == service_j, both counted }
// and afterwards counter[s] has grown by the number of counted calls to s.  That is a stable
// per-service rank; it is computed here without sorting:
//   rank_count_kernel   each CTA owns a contiguous chunk of the batch and histograms it by table
//                       slot in shared memory; the histogram goes to row `blockIdx.x` of a
//                       (chunks x n_slots) matrix
//   rank_scan_kernel    one thread per slot: exclusive scan down its column, counter update
//   rank_assign_kernel  each CTA reloads its row (= calls to each service in earlier chunks) and
//                       walks its chunk 256 calls at a time in order: __match_any_sync ranks the
//                       calls of one service inside a warp, the warps of the CTA take turns on
//                       the running counts, and the seq field of every record is written
// Counters are indexed by table slot; carry_counters_kernel moves them to the slots of the next
// image when the registry flips (services are matched by key).
//
// HBM traffic: two passes over the keys (a 32-byte sector per 48-byte record each) and one 8-byte
// store per record; the (chunks x n_slots) matrix stays in L2.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ptype_b200.h"
#include "device_image.h"
#include "table_hash.h"

namespace ptype {
namespace {

constexpr int kRankThreads = 256;
constexpr uint32_t kNoSlot = 0xFFFFFFFFu;
constexpr uint32_t kFull = 0xFFFFFFFFu;

struct RankTable {
    const ptype_slot* slots;
    uint32_t n_slots;
    uint32_t seed;
};

// Slot of `key` in the cuckoo table, or kNoSlot.  *n_nodes: the service's registered nodes.
__device__ __forceinline__ uint32_t find_slot(const RankTable& t, uint64_t key, uint32_t* n_nodes) {
    const uint32_t lo = static_cast<uint32_t>(key), hi = static_cast<uint32_t>(key >> 32);
#pragma unroll
    for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
        const uint32_t at = slot_of(lo, hi, t.seed, w, t.n_slots);
        const uint4 s = __ldg(reinterpret_cast<const uint4*>(t.slots) + at);
        if (s.z != PTYPE_SLOT_EMPTY && s.x == lo && s.y == hi) {
            *n_nodes = s.w & 0xFFFFFFu;
            return at;
        }
    }
    *n_nodes = 0;
    return kNoSlot;
}

// The slot a call is counted under, or kNoSlot when it consumes no counter value: unknown
// service (no Client exists) or a Client whose list is empty (rpc.go:178-180).
__device__ __forceinline__ uint32_t counted_slot(const RankTable& t, const ptype_req* reqs, size_t j) {
    const uint64_t key = __ldg(reinterpret_cast<const unsigned long long*>(reqs + j));
    uint32_t n_nodes;
    const uint32_t at = find_slot(t, key, &n_nodes);
    return n_nodes ? at : kNoSlot;
}

__global__ void __launch_bounds__(kRankThreads)
    rank_count_kernel(const RankTable t, const ptype_req* __restrict__ reqs, const size_t n, const size_t chunk,
                      uint32_t* __restrict__ rows) {
    extern __shared__ uint32_t hist[];
    for (uint32_t s = threadIdx.x; s < t.n_slots; s += kRankThreads) hist[s] = 0;
    __syncthreads();
    const size_t lo = static_cast<size_t>(blockIdx.x) * chunk;
    const size_t hi = lo + chunk < n ? lo + chunk : n;
    for (size_t j = lo + threadIdx.x; j < hi; j += kRankThreads) {
        const uint32_t at = counted_slot(t, reqs, j);
        if (at != kNoSlot) atomicAdd(&hist[at], 1u);
    }
    __syncthreads();
    uint32_t* row = rows + static_cast<size_t>(blockIdx.x) * t.n_slots;
    for (uint32_t s = threadIdx.x; s < t.n_slots; s += kRankThreads) row[s] = hist[s];
}

// Column s of the matrix becomes its exclusive prefix sum; seq_start[s] keeps the counter value
// the batch started from and the counter moves past the batch.
__global__ void __launch_bounds__(kRankThreads)
    rank_scan_kernel(uint32_t* __restrict__ rows, const int n_chunks, const uint32_t n_slots,
                     unsigned long long* __restrict__ counters, unsigned long long* __restrict__ seq_start) {
    const uint32_t s = blockIdx.x * kRankThreads + threadIdx.x;
    if (s >= n_slots) return;
    uint32_t run = 0;
    for (int c = 0; c < n_chunks; ++c) {
        uint32_t* cell = rows + static_cast<size_t>(c) * n_slots + s;
        const uint32_t v = *cell;
        *cell = run;
        run += v;
    }
    const unsigned long long start = counters[s];
    seq_start[s] = start;
    counters[s] = start + run;
}

__global__ void __launch_bounds__(kRankThreads)
    rank_assign_kernel(const RankTable t, ptype_req* __restrict__ reqs, const size_t n, const size_t chunk,
                       const uint32_t* __restrict__ rows, const unsigned long long* __restrict__ seq_start) {
    extern __shared__ uint32_t run[];  // counted calls to each service before the CTA's cursor
    const uint32_t* row = rows + static_cast<size_t>(blockIdx.x) * t.n_slots;
    for (uint32_t s = threadIdx.x; s < t.n_slots; s += kRankThreads) run[s] = row[s];
    __syncthreads();
    const size_t lo = static_cast<size_t>(blockIdx.x) * chunk;
    const size_t hi = lo + chunk < n ? lo + chunk : n;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (size_t base = lo; base < hi; base += kRankThreads) {  // uniform trip count per CTA
        const size_t j = base + threadIdx.x;
        const uint32_t at = j < hi ? counted_slot(t, reqs, j) : kNoSlot;
        // calls of one service inside this warp: lanes in `peers`, in batch order
        const uint32_t peers = __match_any_sync(kFull, at != kNoSlot ? at : 0x80000000u + lane);
        const uint32_t before = __popc(peers & ((1u << lane) - 1u));
        uint32_t first = 0;
        // the warps hold consecutive calls, so they update the running counts in warp order
        for (uint32_t w = 0; w < kRankThreads / 32; ++w) {
            if (w == warp && at != kNoSlot && before == 0) {
                first = run[at];
                run[at] = first + __popc(peers);
            }
            __syncthreads();
        }
        first = __shfl_sync(kFull, first, __ffs(peers) - 1);
        if (j < hi) {
            const unsigned long long seq = at != kNoSlot ? seq_start[at] + first + before + 1ull : 0ull;
            reinterpret_cast<unsigned long long*>(reqs + j)[1] = seq;  // ptype_req.seq
        }
    }
}

// Counters follow their services from the table of the outgoing image to the table of the
// incoming one (placement may differ; a service the new image does not know is dropped).
__global__ void __launch_bounds__(kRankThreads)
    carry_counters_kernel(const RankTable from, const unsigned long long* __restrict__ from_counters,
                          const RankTable to, unsigned long long* __restrict__ to_counters) {
    const uint32_t s = blockIdx.x * kRankThreads + threadIdx.x;
    if (s >= from.n_slots) return;
    const uint4 e = __ldg(reinterpret_cast<const uint4*>(from.slots) + s);
    if (e.z == PTYPE_SLOT_EMPTY) return;
    uint32_t n_nodes;
    const uint32_t at = find_slot(to, static_cast<uint64_t>(e.x) | (static_cast<uint64_t>(e.y) << 32), &n_nodes);
    if (at != kNoSlot) to_counters[at] = from_counters[s];
}

RankTable table_of(const DeviceImage& img) { return RankTable{img.slots, img.n_slots, img.table_seed}; }

}  // namespace

size_t rank_smem_bytes(uint32_t n_slots) { return static_cast<size_t>(n_slots) * sizeof(uint32_t); }

// CTAs (= chunks) a batch of n calls is split over, given how many CTAs fit the device at once.
cudaError_t rank_plan(uint32_t n_slots, size_t n, int sm_count, int* n_chunks, size_t* chunk) {
    const size_t smem = rank_smem_bytes(n_slots);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(rank_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(rank_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rank_assign_kernel, kRankThreads, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    if (per_sm > 4) per_sm = 4;
    size_t ctas = static_cast<size_t>(sm_count) * per_sm;
    const size_t rounds = (n + kRankThreads - 1) / kRankThreads;
    if (ctas > rounds) ctas = rounds ? rounds : 1;
    // chunk: a whole number of 256-call rounds, so chunk boundaries are warp-aligned
    const size_t per = (rounds + ctas - 1) / ctas;
    *chunk = per * kRankThreads;
    *n_chunks = static_cast<int>((n + *chunk - 1) / *chunk);
    if (*n_chunks < 1) *n_chunks = 1;
    return cudaSuccess;
}

// rows: n_chunks x n_slots uint32 scratch; seq_start: n_slots uint64 scratch.
cudaError_t launch_assign_seq(const DeviceImage& img, ptype_req* reqs, size_t n, int n_chunks, size_t chunk,
                              uint32_t* rows, unsigned long long* counters, unsigned long long* seq_start,
                              cudaStream_t stream) {
    const RankTable t = table_of(img);
    const size_t smem = rank_smem_bytes(img.n_slots);
    rank_count_kernel<<<n_chunks, kRankThreads, smem, stream>>>(t, reqs, n, chunk, rows);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    rank_scan_kernel<<<(img.n_slots + kRankThreads - 1) / kRankThreads, kRankThreads, 0, stream>>>(
        rows, n_chunks, img.n_slots, counters, seq_start);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    rank_assign_kernel<<<n_chunks, kRankThreads, smem, stream>>>(t, reqs, n, chunk, rows, seq_start);
    return cudaGetLastError();
}

cudaError_t launch_carry_counters(const DeviceImage& from, const unsigned long long* from_counters,
                                  const DeviceImage& to, unsigned long long* to_counters, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(to_counters, 0, static_cast<size_t>(to.n_slots) * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    carry_counters_kernel<<<(from.n_slots + kRankThreads - 1) / kRankThreads, kRankThreads, 0, stream>>>(
        table_of(from), from_counters, table_of(to), to_counters);
    return cudaGetLastError();
}

}  // namespace ptype
