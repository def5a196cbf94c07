// device_image.h — what a dispatch kernel is handed by value: pointers into the
// device copy of the registry image plus the per-context limits.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ptype_b200.h"

namespace ptype {

struct DeviceImage {
    const ptype_slot* slots;     // n_slots entries, 16-byte aligned
    const uint32_t* conn_hash;   // max_conn entries: FNV-1a-32(localAddr+itoa(i))
    const ptype_method* methods; // n_methods entries (error paths)
    const ptype_method* mtable;  // PTYPE_MTABLE_SLOTS entries, perfect hash
    uint32_t n_slots;
    uint32_t max_conn;           // ConnConfig.MaxConnections, 0 = mesh
    uint32_t n_methods;
    uint32_t table_seed;         // cuckoo hash seed
    uint32_t method_mul;         // perfect-hash multiplier
    uint32_t check_only;         // 1: the only receiver any service exposes is Prime (ALU-heavy calls)
    uint64_t m_magic;            // 2^64 / max_conn + 1: multiply-shift modulo by max_conn
    uint64_t check_limit;        // Prime.Check trip-count limit
};

// Layout of the device-side ptype_batch_stats accumulator (uint64 words).
enum : int { kStatCalls = 0, kStatByStatus = 1, kStatFailed = 1 + PTYPE_ST__COUNT, kStatWords = 2 + PTYPE_ST__COUNT };

struct LaunchPlan {
    int threads = 0;        // per CTA
    int ctas_per_sm = 0;    // resident CTAs per SM (occupancy query)
    size_t dyn_smem = 0;    // dynamic shared memory per CTA: staged table + request stages
    size_t table_smem = 0;  // bytes of it holding the slot table (0 = probe in L2)
    bool smem_table = false;
    bool bulk = false;      // requests staged by bulk-async copies (else LDG + shuffle)
};

// Number of partial sums the checksum kernel writes to its scratch buffer.
constexpr int kChecksumBlocks = 148 * 8;
constexpr int kChecksumThreads = 256;

// Must run with the target device current (occupancy query of the kernel the plan names).
// `pdl` in the launchers below: the launch may overlap the tail of the previous kernel on the
// stream (programmatic dependent launch); pass true only when that kernel was a dispatch launch.
cudaError_t plan_dispatch(const DeviceImage& img, uint32_t flags, bool extras, LaunchPlan* plan);
int grid_for(const LaunchPlan& plan, size_t n, int sm_count);
cudaError_t launch_dispatch(const DeviceImage& img, const LaunchPlan& plan, int blocks,
                            const ptype_req* reqs, size_t n, ptype_reply* replies, uint32_t* failed_idx,
                            uint32_t failed_cap, uint32_t idx_base, unsigned long long* stats,
                            cudaStream_t stream, bool pdl);
int max_batches_per_launch();
// seq != nullptr (one batch, not a bulk plan): call j takes seq[j] instead of its record's seq field.
cudaError_t launch_dispatch_multi(const DeviceImage& img, const LaunchPlan& plan, int sm_count,
                                  const ptype_batch* batches, int count, cudaStream_t stream, bool pdl,
                                  const unsigned long long* seq = nullptr, int* kernels = nullptr,
                                  bool lone = false);
// Values-only output of a client batch: values != nullptr switches the kernel from 16-byte
// reply records to 8-byte values plus an (index, status) list of the failed calls.
struct CallBatchOut {
    int64_t* values = nullptr;
    ptype_failed* failed = nullptr;
    uint32_t failed_cap = 0;
    uint32_t index_base = 0;
    unsigned long long* failed_count = nullptr;
    // Single-launch batches only: the last CTA to finish copies the failure count to
    // *host_count (mapped page-locked memory) and zeroes failed_count and done_ctas again, so
    // the caller needs no memset before the launch and no copy after it.
    unsigned long long* done_ctas = nullptr;
    unsigned long long* host_count = nullptr;
};
cudaError_t launch_call_batch(const DeviceImage& img, const LaunchPlan& plan, int blocks,
                              uint64_t svc_key, uint64_t seq_base, const uint32_t* methods,
                              const uint32_t* recvs, int n_methods, const int64_t* args,
                              int args_per_call, const uint8_t* method_sel, size_t n,
                              ptype_reply* replies, const CallBatchOut& out, cudaStream_t stream, bool pdl);
cudaError_t launch_reply_checksum(const ptype_reply* replies, size_t n, uint64_t index_base,
                                  unsigned long long* partials, unsigned long long* out,
                                  cudaStream_t stream);

cudaError_t launch_resolve_names(const unsigned char* names, uint32_t stride, size_t n, ptype_method_ref* out,
                                 ptype_req* patch, cudaStream_t stream);
size_t gob_scratch_bytes(size_t n);
cudaError_t launch_gob_encode_replies(const ptype_reply* replies, size_t n, unsigned long long* offsets,
                                      uint8_t* out, unsigned long long out_cap, void* scratch,
                                      unsigned long long* total, cudaStream_t stream);
cudaError_t launch_gob_decode_args(const uint8_t* bytes, const unsigned long long* offsets, size_t n,
                                   int64_t* args, int args_per_call, uint8_t* status, cudaStream_t stream);

// wire.cu: net/rpc byte streams.  call_off / out_off have n_streams + 1 entries.
cudaError_t launch_wire_count(const uint8_t* bytes, const unsigned long long* stream_off, size_t n_streams,
                              unsigned long long* call_off, cudaStream_t stream);
cudaError_t launch_wire_decode(const uint8_t* bytes, const unsigned long long* stream_off, size_t n_streams,
                               const ptype_wire_stream* streams, unsigned long long* call_off, ptype_req* reqs,
                               ptype_wire_call* calls, cudaStream_t stream);
cudaError_t launch_wire_encode(bool size_only, const uint8_t* req_bytes, const unsigned long long* call_off,
                               size_t n_streams, const ptype_wire_call* calls, const ptype_reply* replies,
                               const uint8_t* defs, uint32_t response_def_len, uint32_t empty_def_len,
                               uint32_t response_id, uint32_t empty_id, uint8_t* state, unsigned long long* out_off,
                               uint8_t* out, cudaStream_t stream);

cudaError_t launch_patch_service(ptype_slot* slots, const uint32_t* order, uint32_t n_slots, uint32_t seed,
                                 uint64_t key, int delta, uint32_t* result, cudaStream_t stream);

// rank.cu: balancer counters on the device.
struct RankPlan {
    bool staged = false;  // count pass: table keys staged in shared memory (else probed in L1/L2)
    int threads = 0;      // count pass: threads per CTA
    size_t smem = 0;      // count pass: shared memory per CTA
    int count_ctas = 0;   // count pass: CTAs, each taking `sub` consecutive chunks
    int sub = 1;
    int n_chunks = 0;     // contiguous chunks of the batch = rows of the histogram matrix = assign CTAs
    size_t chunk = 0;     // records per chunk
    size_t assign_smem = 0;
};
bool rank_supported(uint32_t n_slots);
cudaError_t rank_plan(uint32_t n_slots, size_t n, int sm_count, RankPlan* plan);
cudaError_t launch_assign_seq(const DeviceImage& img, const RankPlan& plan, ptype_req* reqs, size_t n,
                              uint32_t* rows, uint32_t* ids, unsigned long long* counters,
                              unsigned long long* seq_start, unsigned long long* seq_out, cudaStream_t stream);
cudaError_t launch_carry_counters(const DeviceImage& from, const unsigned long long* from_counters,
                                  const DeviceImage& to, unsigned long long* to_counters, cudaStream_t stream);

}  // namespace ptype
#include <string>
namespace ptype {
// The gob type-definition messages of rpc.Response and of the empty struct for the given type ids.
void wire_server_typedefs(uint32_t response_id, uint32_t empty_id, std::string* response_def, std::string* empty_def);

// Records the text ptype_b200_last_error() returns on the calling thread.
void set_last_error(const char* text);

}  // namespace ptype
