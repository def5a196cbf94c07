// rank.cu — the balancer's counters on the device.
//
// Every Client of the reference owns one counter; each call that finds a non-empty client list
// takes its next value (cluster/rpc.go:129,172-183: `if len(clients) == 0 { return nil }` comes
// BEFORE `atomic.AddUint64(&c.seq, 1)`, so calls to a service without nodes consume nothing).
// A batch that mixes services is N such calls in batch order, so call j gets
//     seq_j = counter[service_j] + 1 + #{ i < j : service_i == service_j, both counted }
// and afterwards counter[s] has grown by the number of counted calls to s.  That is a stable
// per-service rank; it is computed here without sorting:
//   rank_count_kernel   the batch is cut into contiguous chunks; a CTA histograms a chunk by table
//                       slot in shared memory, stores the histogram as row `chunk` of a
//                       (chunks x n_slots) matrix, and leaves behind the slot id of every record
//                       (4 bytes per record), so that nobody has to look a key up again
//   rank_scan_kernel    exclusive scan down every column (eight warps per 32 columns), counter update
//   rank_assign_kernel  one 256-thread CTA per chunk reloads its row (= calls to each service in
//                       earlier chunks) and walks the chunk's slot ids in order, each warp holding
//                       256 consecutive calls per trip: ballots over the slot bits rank the calls of one
//                       service inside a tile of 32, the warps of the CTA take turns on the running
//                       counts (one turn settles a warp's eight tiles; only the tile leaders touch
//                       shared memory inside a turn), and the seq of every record is written
// Counters are indexed by table slot; carry_counters_kernel moves them to the slots of the next
// image when the registry flips (services are matched by key).
//
// The count pass stages the table's keys in shared memory where they fit (tables up to ~18 500
// slots, 12 bytes per slot, one 1024-thread CTA per SM taking several chunks in turn): three random
// 16-byte probes per record against a table in L2 cost more than everything else together
// (measured on the 10 000-service registry: 497 us per 2^24 records).  Larger tables, up to 51 200
// slots, are probed in L1/L2 by 256-thread CTAs.  The assign pass needs 4 bytes of shared memory
// per slot, so four of its CTAs share an SM on that registry and their turn chains interleave.
//
// HBM traffic: the count pass reads the keys (an 8-byte key drags its 48-byte record along: DRAM
// moves 64 bytes at a time) and writes 4 bytes per record; the assign pass reads those 4 bytes and
// stores 8 bytes into every 48-byte record, which the memory system turns into a read and a write
// of the 32-byte sector (tools/partial_write_probe.cu: 522 MB read + 518 MB written for 2^24
// records, 184 us on its own) - that store, not the ranking, is the floor of this pass.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ptype_b200.h"
#include "device_image.h"
#include "table_hash.h"

namespace ptype {
namespace {

constexpr int kStagedThreads = 1024;  // keys in shared memory: one CTA per SM (two for small tables)
constexpr int kProbeThreads = 256;    // table probed in L1/L2: up to four CTAs per SM
constexpr int kScanThreads = 256;
constexpr uint32_t kNoSlot = 0xFFFFFFFFu;
constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr size_t kRankSmemBudget = 220 * 1024;

struct RankTable {
    const ptype_slot* slots;
    uint32_t n_slots;
    uint32_t seed;
};

// Shared memory of a rank CTA: [keys: 8 B per slot][counts: 4 B per slot][countable: 1 bit per
// slot]; the first and last part only when the keys are staged.
template <bool kStaged>
struct RankShared {
    const unsigned long long* keys;
    uint32_t* counts;
    const uint32_t* countable;
};

__host__ __device__ inline size_t rank_shared_bytes(bool staged, uint32_t n_slots) {
    return staged ? static_cast<size_t>(n_slots) * 12 + static_cast<size_t>((n_slots + 31) / 32) * 4
                  : static_cast<size_t>(n_slots) * 4;
}

__device__ __forceinline__ uint4 ld_slot(const ptype_slot* slot) {
    return __ldg(reinterpret_cast<const uint4*>(slot));
}

// Carves the CTA's dynamic shared memory and, when staging, copies the table's keys in and
// marks the slots that count: occupied, and the service has at least one node.
template <bool kStaged>
__device__ __forceinline__ RankShared<kStaged> rank_shared(const RankTable& t, unsigned char* smem) {
    RankShared<kStaged> sh;
    if (!kStaged) {
        sh.keys = nullptr;
        sh.countable = nullptr;
        sh.counts = reinterpret_cast<uint32_t*>(smem);
        return sh;
    }
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem);
    sh.counts = reinterpret_cast<uint32_t*>(smem + static_cast<size_t>(t.n_slots) * 8);
    uint32_t* countable = sh.counts + t.n_slots;
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t base = (threadIdx.x >> 5) * 32u; base < t.n_slots; base += blockDim.x) {
        const uint32_t s = base + lane;
        bool counts_here = false;
        if (s < t.n_slots) {
            const uint4 e = ld_slot(t.slots + s);
            keys[s] = static_cast<unsigned long long>(e.x) | (static_cast<unsigned long long>(e.y) << 32);
            counts_here = e.z != PTYPE_SLOT_EMPTY && (e.w & 0xFFFFFFu) != 0;
        }
        const uint32_t bits = __ballot_sync(kFull, counts_here);
        if (lane == 0) countable[base >> 5] = bits;
    }
    sh.keys = keys;
    sh.countable = countable;
    return sh;
}

// The slot a call is counted under, or kNoSlot when it consumes no counter value: unknown
// service (no Client exists) or a Client whose list is empty (rpc.go:178-180).  All three
// candidates are read unconditionally: independent loads, one latency.
template <bool kStaged>
__device__ __forceinline__ uint32_t counted_slot(const RankTable& t, const RankShared<kStaged>& sh,
                                                 unsigned long long key) {
    const uint32_t lo = static_cast<uint32_t>(key), hi = static_cast<uint32_t>(key >> 32);
    uint32_t at[PTYPE_SLOT_WAYS];
    bool hit[PTYPE_SLOT_WAYS];
    if (kStaged) {
        unsigned long long k[PTYPE_SLOT_WAYS];
        uint32_t bits[PTYPE_SLOT_WAYS];
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
            at[w] = slot_of(lo, hi, t.seed, w, t.n_slots);
            k[w] = sh.keys[at[w]];
            bits[w] = sh.countable[at[w] >> 5];
        }
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) hit[w] = k[w] == key && ((bits[w] >> (at[w] & 31u)) & 1u);
    } else {
        uint4 e[PTYPE_SLOT_WAYS];
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
            at[w] = slot_of(lo, hi, t.seed, w, t.n_slots);
            e[w] = ld_slot(t.slots + at[w]);
        }
#pragma unroll
        for (int w = 0; w < PTYPE_SLOT_WAYS; ++w)
            hit[w] = e[w].z != PTYPE_SLOT_EMPTY && e[w].x == lo && e[w].y == hi && (e[w].w & 0xFFFFFFu) != 0;
    }
    uint32_t found = kNoSlot;
#pragma unroll
    for (int w = PTYPE_SLOT_WAYS - 1; w >= 0; --w)
        if (hit[w]) found = at[w];
    return found;
}

// Records handled per thread and loop trip.  Their keys are loaded one trip ahead and their
// probes are independent of one another.  A warp's records are CONSECUTIVE: kRankDepth tiles of
// 32, so that in rank_assign_kernel a warp settles all of them in one turn.
constexpr int kRankDepth = 4;

// Index of the r-th record of this thread in the trip that starts at `base`.
__device__ __forceinline__ size_t record_index(size_t base, int r) {
    return base + (static_cast<size_t>(threadIdx.x >> 5) * kRankDepth + r) * 32 + (threadIdx.x & 31u);
}

// Keys of this thread's records; addresses past `hi` are clamped.
__device__ __forceinline__ void load_keys(const ptype_req* reqs, size_t base, size_t hi,
                                          unsigned long long (&key)[kRankDepth]) {
#pragma unroll
    for (int r = 0; r < kRankDepth; ++r) {
        const size_t j = record_index(base, r);
        key[r] = __ldg(reinterpret_cast<const unsigned long long*>(reqs + (j < hi ? j : hi - 1)));
    }
}

template <bool kStaged>
__device__ __forceinline__ void counted_slots(const RankTable& t, const RankShared<kStaged>& sh,
                                              const unsigned long long (&key)[kRankDepth], size_t base, size_t hi,
                                              uint32_t (&at)[kRankDepth]) {
#pragma unroll
    for (int r = 0; r < kRankDepth; ++r) {
        const size_t j = record_index(base, r);
        const uint32_t slot = counted_slot<kStaged>(t, sh, key[r]);
        at[r] = j < hi ? slot : kNoSlot;
    }
}

// One CTA histograms `sub` consecutive chunks (rows of the matrix), one after the other, and leaves
// behind the slot every record is counted under (ids[j], kNoSlot for a record that consumes no
// counter value): the assign pass then needs neither the keys nor the table.
template <bool kStaged, int kThreads>
__global__ void __launch_bounds__(kThreads)
    rank_count_kernel(const RankTable t, const ptype_req* __restrict__ reqs, const size_t n, const size_t chunk,
                      const int sub, const int n_chunks, uint32_t* __restrict__ rows, uint32_t* __restrict__ ids) {
    extern __shared__ __align__(16) unsigned char rank_smem[];
    const RankShared<kStaged> sh = rank_shared<kStaged>(t, rank_smem);
    const uint32_t lane = threadIdx.x & 31u;
    constexpr size_t kTrip = static_cast<size_t>(kThreads) * kRankDepth;
    for (int k = 0; k < sub; ++k) {
        const int c = static_cast<int>(blockIdx.x) * sub + k;
        if (c >= n_chunks) break;
        for (uint32_t s = threadIdx.x; s < t.n_slots; s += kThreads) sh.counts[s] = 0;
        __syncthreads();
        const size_t lo = static_cast<size_t>(c) * chunk;
        const size_t hi = lo + chunk < n ? lo + chunk : n;
        unsigned long long key[kRankDepth], next[kRankDepth];
        if (lo < hi) load_keys(reqs, lo, hi, next);
        for (size_t base = lo; base < hi; base += kTrip) {
#pragma unroll
            for (int r = 0; r < kRankDepth; ++r) key[r] = next[r];
            if (base + kTrip < hi) load_keys(reqs, base + kTrip, hi, next);
            uint32_t at[kRankDepth];
            counted_slots<kStaged>(t, sh, key, base, hi, at);
#pragma unroll
            for (int r = 0; r < kRankDepth; ++r) {
                const size_t j = record_index(base, r);
                if (j < hi) ids[j] = at[r];
                // A tile that names one service only (a single Client's batch) adds once; otherwise
                // every counted record adds for itself - __match_any_sync would merge equal slots, but
                // its latency was ~45 % of this kernel's stall samples on the 10 000-service stream.
                const uint32_t lead = __shfl_sync(kFull, at[r], 0);
                if (__all_sync(kFull, at[r] == lead)) {
                    if (lane == 0 && lead != kNoSlot) atomicAdd(&sh.counts[lead], 32u);
                } else if (at[r] != kNoSlot) {
                    atomicAdd(&sh.counts[at[r]], 1u);
                }
            }
        }
        __syncthreads();
        uint32_t* row = rows + static_cast<size_t>(c) * t.n_slots;
        for (uint32_t s = threadIdx.x; s < t.n_slots; s += kThreads) row[s] = sh.counts[s];
        __syncthreads();
    }
}

// Column s of the matrix becomes its exclusive prefix sum; seq_start[s] keeps the counter value
// the batch started from and the counter moves past the batch.  A CTA takes 32 adjacent columns
// (a row segment is one 128-byte line); its eight warps split the rows into eight runs, sum
// their runs, exchange the sums, then rewrite their runs - a column's latency chain is a
// sixteenth of walking it top to bottom, and the loads of eight rows are always in flight.
constexpr int kScanRows = 8;
__global__ void __launch_bounds__(kScanThreads)
    rank_scan_kernel(uint32_t* __restrict__ rows, const int n_chunks, const uint32_t n_slots,
                     unsigned long long* __restrict__ counters, unsigned long long* __restrict__ seq_start) {
    __shared__ uint32_t warp_sum[kScanThreads / 32][32];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t s = blockIdx.x * 32u + lane;
    const bool live = s < n_slots;
    constexpr int kWarps = kScanThreads / 32;
    const int per = (n_chunks + kWarps - 1) / kWarps;
    const int c0 = static_cast<int>(warp) * per < n_chunks ? static_cast<int>(warp) * per : n_chunks;
    const int c1 = c0 + per < n_chunks ? c0 + per : n_chunks;
    uint32_t* col = rows + (live ? s : 0);
    uint32_t sum = 0;
    if (live) {
        int c = c0;
        for (; c + kScanRows <= c1; c += kScanRows) {
            uint32_t v[kScanRows];
#pragma unroll
            for (int k = 0; k < kScanRows; ++k) v[k] = col[static_cast<size_t>(c + k) * n_slots];
#pragma unroll
            for (int k = 0; k < kScanRows; ++k) sum += v[k];
        }
        for (; c < c1; ++c) sum += col[static_cast<size_t>(c) * n_slots];
    }
    warp_sum[warp][lane] = sum;
    __syncthreads();
    uint32_t run = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
        const uint32_t v = warp_sum[w][lane];
        if (w < static_cast<int>(warp)) run += v;
        total += v;
    }
    if (!live) return;
    int c = c0;
    for (; c + kScanRows <= c1; c += kScanRows) {
        uint32_t v[kScanRows];
#pragma unroll
        for (int k = 0; k < kScanRows; ++k) v[k] = col[static_cast<size_t>(c + k) * n_slots];
#pragma unroll
        for (int k = 0; k < kScanRows; ++k) {
            col[static_cast<size_t>(c + k) * n_slots] = run;
            run += v[k];
        }
    }
    for (; c < c1; ++c) {
        const uint32_t v = col[static_cast<size_t>(c) * n_slots];
        col[static_cast<size_t>(c) * n_slots] = run;
        run += v;
    }
    if (warp == 0) {
        const unsigned long long start = counters[s];
        seq_start[s] = start;
        counters[s] = start + total;
    }
}

// One 256-thread CTA per chunk (row).  It loads the row (calls to each service in earlier chunks)
// into shared memory and walks the chunk's slot ids in order: a warp holds kAssignDepth consecutive
// tiles of 32 records per trip, ballots over the slot bits rank the records of one service inside a tile,
// and the warps take turns on the running counts - a turn settles all of a warp's tiles.  Only the
// leader lane of each service in a tile touches shared memory inside the turn; everything else
// (loading ids, matching, fetching the counter base, storing) happens outside it.  Several CTAs
// share an SM (the counts are all the shared memory a CTA needs), so their turn chains interleave.
// kScatter: write ptype_req.seq of the records; else a plain array seq_out[j].
// Tried instead of the turns: the eight warps each own a slot class (slot & 7), every trip's records
// are sorted into eight class lists in shared memory and each warp settles its own list, nobody
// waiting for anybody.  Correct, and slower (394 vs 338 us per 2^24 records for the whole
// assign_seq): the sorting and the per-entry bookkeeping cost five times the instructions
// (150 M vs 31 M warp instructions) and the pass became issue-bound instead of latency-bound.
constexpr int kAssignThreads = 256;
constexpr int kAssignDepth = 8;
// Lanes of the warp that name the same slot: one ballot per bit of the slot id.  With thousands of
// services nearly all 32 values of a tile differ, the case in which MATCH.ANY is slowest; the
// ballots took the whole assign_seq from 368 to 338 us per 2^24 records.
__device__ __forceinline__ uint32_t match_slots(uint32_t slot, uint32_t lane, uint32_t bits) {
    const bool live = slot != kNoSlot;
    uint32_t differ = 0;  // lanes whose slot differs from this lane's in some bit
    for (uint32_t b = 0; b < bits; ++b) {
        const bool one = (slot >> b) & 1u;
        differ |= __ballot_sync(kFull, one) ^ (one ? 0xFFFFFFFFu : 0u);
    }
    const uint32_t peers = __ballot_sync(kFull, live) & ~differ;
    return live ? peers : 1u << lane;
}
template <bool kScatter>
__global__ void __launch_bounds__(kAssignThreads)
    rank_assign_kernel(const uint32_t n_slots, const uint32_t* __restrict__ ids, ptype_req* __restrict__ reqs,
                       unsigned long long* __restrict__ seq_out, const size_t n, const size_t chunk,
                       const uint32_t* __restrict__ rows, const unsigned long long* __restrict__ seq_start) {
    extern __shared__ __align__(16) unsigned char rank_smem[];
    uint32_t* run = reinterpret_cast<uint32_t*>(rank_smem);  // counted calls to each service before the cursor
    const uint32_t* row = rows + static_cast<size_t>(blockIdx.x) * n_slots;
    for (uint32_t s = threadIdx.x; s < n_slots; s += kAssignThreads) run[s] = row[s];
    __syncthreads();
    const size_t lo = static_cast<size_t>(blockIdx.x) * chunk;
    const size_t hi = lo + chunk < n ? lo + chunk : n;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t slot_bits = 32u - __clz(n_slots - 1u);
    constexpr size_t kTrip = static_cast<size_t>(kAssignThreads) * kAssignDepth;
    const size_t lane_off = static_cast<size_t>(warp) * kAssignDepth * 32 + lane;
    for (size_t base = lo; base < hi; base += kTrip) {  // uniform trip count per CTA
        const size_t mine = base + lane_off;
        uint32_t at[kAssignDepth], peers[kAssignDepth], first[kAssignDepth];
        unsigned long long start[kAssignDepth];  // the counters' values before the batch: fetched now,
                                                 // needed after the turns, so their latency hides there
#pragma unroll
        for (int r = 0; r < kAssignDepth; ++r) {
            const size_t j = mine + static_cast<size_t>(r) * 32;
            at[r] = j < hi ? __ldg(ids + j) : kNoSlot;
        }
#pragma unroll
        for (int r = 0; r < kAssignDepth; ++r) start[r] = at[r] != kNoSlot ? __ldg(seq_start + at[r]) : 0ull;
#pragma unroll
        for (int r = 0; r < kAssignDepth; ++r) {
            // calls of one service inside a tile of 32: lanes in `peers`, in batch order
            peers[r] = match_slots(at[r], lane, slot_bits);
            first[r] = 0;
        }
        for (uint32_t w = 0; w < kAssignThreads / 32; ++w) {
            if (w == warp) {
                // Only a tile's leaders (first lane of each service in the tile: distinct slots,
                // so no two lanes of an instruction meet) touch the running counts; tiles in order,
                // because the next tile may name the same service from another lane.  (A shared-
                // memory atomic per tile instead of the load/store pair measured slower: 286 vs
                // 220 us per 2^24 records - 32 spread lanes cost an atomic 64 LSU cycles.)
#pragma unroll
                for (int r = 0; r < kAssignDepth; ++r) {
                    if (at[r] != kNoSlot && (peers[r] & ((1u << lane) - 1u)) == 0) {
                        first[r] = run[at[r]];
                        run[at[r]] = first[r] + __popc(peers[r]);
                    }
                    __syncwarp();
                }
            }
            __syncthreads();
        }
#pragma unroll
        for (int r = 0; r < kAssignDepth; ++r) {
            const size_t j = mine + static_cast<size_t>(r) * 32;
            const uint32_t f = __shfl_sync(kFull, first[r], __ffs(peers[r]) - 1);
            if (j < hi) {
                unsigned long long seq = 0ull;
                if (at[r] != kNoSlot) seq = start[r] + f + __popc(peers[r] & ((1u << lane) - 1u)) + 1ull;
                if (kScatter)
                    reinterpret_cast<unsigned long long*>(reqs + j)[1] = seq;  // ptype_req.seq
                else
                    seq_out[j] = seq;
            }
        }
    }
}

// Counters follow their services from the table of the outgoing image to the table of the
// incoming one (placement may differ; a service the new image does not know is dropped).
__global__ void __launch_bounds__(kScanThreads)
    carry_counters_kernel(const RankTable from, const unsigned long long* __restrict__ from_counters,
                          const RankTable to, unsigned long long* __restrict__ to_counters) {
    const uint32_t s = blockIdx.x * kScanThreads + threadIdx.x;
    if (s >= from.n_slots) return;
    const uint4 e = ld_slot(from.slots + s);
    if (e.z == PTYPE_SLOT_EMPTY) return;
    const uint32_t lo = e.x, hi = e.y;
#pragma unroll
    for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
        const uint32_t at = slot_of(lo, hi, to.seed, w, to.n_slots);
        const uint4 d = ld_slot(to.slots + at);
        if (d.z != PTYPE_SLOT_EMPTY && d.x == lo && d.y == hi) {
            to_counters[at] = from_counters[s];
            return;
        }
    }
}

RankTable table_of(const DeviceImage& img) { return RankTable{img.slots, img.n_slots, img.table_seed}; }

template <bool kStaged, int kThreads>
cudaError_t prepare_count(size_t smem) {
    if (smem <= 32 * 1024) return cudaSuccess;  // well below the 48 KiB default: asking for less is harmless
    return cudaFuncSetAttribute(rank_count_kernel<kStaged, kThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(smem));
}

}  // namespace

bool rank_supported(uint32_t n_slots) { return rank_shared_bytes(false, n_slots) <= 200 * 1024; }

// How a batch of n calls is split.  Rows of the matrix = chunks = CTAs of the assign pass (256
// threads, as many per SM as their counts fit); a CTA of the count pass (which may stage the keys
// and then fits once per SM) takes `sub` consecutive chunks.
cudaError_t rank_plan(uint32_t n_slots, size_t n, int sm_count, RankPlan* plan) {
    plan->staged = rank_shared_bytes(true, n_slots) <= kRankSmemBudget;
    plan->threads = plan->staged ? kStagedThreads : kProbeThreads;
    plan->smem = rank_shared_bytes(plan->staged, n_slots);
    plan->assign_smem = static_cast<size_t>(n_slots) * 4;
    cudaError_t e = plan->staged ? prepare_count<true, kStagedThreads>(plan->smem)
                                 : prepare_count<false, kProbeThreads>(plan->smem);
    if (e != cudaSuccess) return e;
    int assign_per_sm = 0;
    for (int scatter = 0; scatter < 2; ++scatter) {
        auto kernel = scatter ? rank_assign_kernel<true> : rank_assign_kernel<false>;
        if (plan->assign_smem > 32 * 1024) {
            e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(plan->assign_smem));
            if (e != cudaSuccess) return e;
        }
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&assign_per_sm, kernel, kAssignThreads, plan->assign_smem);
        if (e != cudaSuccess) return e;
    }
    if (assign_per_sm < 1) return cudaErrorInvalidConfiguration;
    if (assign_per_sm > 8) assign_per_sm = 8;
    // a chunk is a whole number of the assign pass's trips, so chunk boundaries are warp-aligned
    const size_t trip = static_cast<size_t>(kAssignThreads) * kAssignDepth;
    const size_t trips = (n + trip - 1) / trip;
    size_t rows = static_cast<size_t>(sm_count) * assign_per_sm;
    if (rows > trips) rows = trips ? trips : 1;
    plan->chunk = (trips + rows - 1) / rows * trip;
    plan->n_chunks = static_cast<int>((n + plan->chunk - 1) / plan->chunk);
    if (plan->n_chunks < 1) plan->n_chunks = 1;
    // count pass: one CTA per SM when it stages the keys (two for small tables), four when it probes
    int count_ctas = sm_count * (plan->staged ? (plan->smem <= 100 * 1024 ? 2 : 1) : 4);
    if (count_ctas > plan->n_chunks) count_ctas = plan->n_chunks;
    plan->sub = (plan->n_chunks + count_ctas - 1) / count_ctas;
    plan->count_ctas = (plan->n_chunks + plan->sub - 1) / plan->sub;
    return cudaSuccess;
}

// rows: n_chunks x n_slots uint32 scratch; ids: n uint32 scratch; seq_start: n_slots uint64 scratch.
// seq_out == nullptr: the sequence numbers go into ptype_req.seq of the records; else into seq_out[j].
cudaError_t launch_assign_seq(const DeviceImage& img, const RankPlan& plan, ptype_req* reqs, size_t n,
                              uint32_t* rows, uint32_t* ids, unsigned long long* counters,
                              unsigned long long* seq_start, unsigned long long* seq_out, cudaStream_t stream) {
    const RankTable t = table_of(img);
    if (plan.staged)
        rank_count_kernel<true, kStagedThreads><<<plan.count_ctas, kStagedThreads, plan.smem, stream>>>(
            t, reqs, n, plan.chunk, plan.sub, plan.n_chunks, rows, ids);
    else
        rank_count_kernel<false, kProbeThreads><<<plan.count_ctas, kProbeThreads, plan.smem, stream>>>(
            t, reqs, n, plan.chunk, plan.sub, plan.n_chunks, rows, ids);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    rank_scan_kernel<<<(img.n_slots + 31) / 32, kScanThreads, 0, stream>>>(rows, plan.n_chunks, img.n_slots, counters,
                                                                          seq_start);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (seq_out == nullptr)
        rank_assign_kernel<true><<<plan.n_chunks, kAssignThreads, plan.assign_smem, stream>>>(
            img.n_slots, ids, reqs, nullptr, n, plan.chunk, rows, seq_start);
    else
        rank_assign_kernel<false><<<plan.n_chunks, kAssignThreads, plan.assign_smem, stream>>>(
            img.n_slots, ids, nullptr, seq_out, n, plan.chunk, rows, seq_start);
    return cudaGetLastError();
}

cudaError_t launch_carry_counters(const DeviceImage& from, const unsigned long long* from_counters,
                                  const DeviceImage& to, unsigned long long* to_counters, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(to_counters, 0, static_cast<size_t>(to.n_slots) * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    carry_counters_kernel<<<(from.n_slots + kScanThreads - 1) / kScanThreads, kScanThreads, 0, stream>>>(
        table_of(from), from_counters, table_of(to), to_counters);
    return cudaGetLastError();
}

}  // namespace ptype
