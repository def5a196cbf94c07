// context.cu — the C ABI over the kernels: contexts, snapshot installation,
// host- and device-pointer dispatch, NCCL broadcast of registry images.
//
// No CPU fallback exists: every dispatch entry point launches the sm_100a
// kernels or returns an error.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/ptype_b200.h"
#include "device_image.h"
#include "table_hash.h"

namespace {

thread_local std::string g_last_error;

int cuda_fail(cudaError_t e, const char* what) {
    char buf[512];
    std::snprintf(buf, sizeof buf, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    g_last_error = buf;
    return PTYPE_E_CUDA;
}

#define CU(call)                                         \
    do {                                                 \
        cudaError_t e_ = (call);                         \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call); \
    } while (0)

constexpr int kPipe = 4;                     // host-pointer pipeline depth
constexpr size_t kMaxChunkCalls = 1024 * 1024;  // staging capacity per pipeline slot
// Calls per pipelined chunk of the host-buffer paths.  PTYPE_B200_CHUNK overrides it
// (measurement switch; multiples of 32, at most kMaxChunkCalls).
size_t chunk_calls() {
    static const size_t v = [] {
        size_t c = 128 * 1024;
        if (const char* e = std::getenv("PTYPE_B200_CHUNK")) {
            const long long x = std::atoll(e);
            if (x >= 32 && static_cast<size_t>(x) <= kMaxChunkCalls) c = static_cast<size_t>(x) / 32 * 32;
        }
        return c;
    }();
    return v;
}

// ---- NCCL, resolved at run time so the library has no link-time NCCL dependency
// (inside a torch process this picks up the libnccl torch already loaded).
struct UniqueId128 {
    char internal[128];
};
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, UniqueId128, int) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclUint8 = 1;  // ncclUint8 in nccl.h's ncclDataType_t

NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) return;
        api.GetUniqueId = reinterpret_cast<int (*)(void*)>(dlsym(api.handle, "ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<int (*)(void**, int, UniqueId128, int)>(
            dlsym(api.handle, "ncclCommInitRank"));
        api.Broadcast =
            reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(
                dlsym(api.handle, "ncclBroadcast"));
        api.CommDestroy = reinterpret_cast<int (*)(void*)>(dlsym(api.handle, "ncclCommDestroy"));
        api.GetErrorString =
            reinterpret_cast<const char* (*)(int)>(dlsym(api.handle, "ncclGetErrorString"));
    });
    if (!api.handle || !api.GetUniqueId || !api.CommInitRank || !api.Broadcast) return nullptr;
    return &api;
}

int nccl_fail(int r, const char* what) {
    NcclApi* a = nccl_api();
    char buf[512];
    std::snprintf(buf, sizeof buf, "%s: %s", what,
                  (a && a->GetErrorString) ? a->GetErrorString(r) : "nccl error");
    g_last_error = buf;
    return PTYPE_E_NCCL;
}

// Launch-plan variants selectable through the dispatch flags (measurement switches).
constexpr int kPlanKinds = 6;
constexpr uint32_t kPlanFlags[kPlanKinds] = {0u,
                                             PTYPE_F_GLOBAL_TABLE,
                                             PTYPE_F_LDG_LOADS,
                                             PTYPE_F_GLOBAL_TABLE | PTYPE_F_LDG_LOADS,
                                             PTYPE_F_BULK_LOADS,
                                             PTYPE_F_GLOBAL_TABLE | PTYPE_F_BULK_LOADS};
inline int plan_kind(uint32_t flags) {
    const int load = (flags & PTYPE_F_LDG_LOADS) ? 1 : ((flags & PTYPE_F_BULK_LOADS) ? 2 : 0);
    return load * 2 + ((flags & PTYPE_F_GLOBAL_TABLE) ? 1 : 0);
}

struct Stage {
    ptype_req* d_reqs = nullptr;
    ptype_reply* d_replies = nullptr;
    cudaEvent_t done = nullptr;
    bool busy = false;
};

struct DeviceState {
    int device = -1;
    int sm_count = 0;
    int dispatch_sms = 0;  // SMs the dispatch grids are sized for (see ptype_b200_comm_init)
    cudaStream_t stream = nullptr;       // device-pointer dispatches, checksums
    // True while the last thing enqueued on `stream` is a dispatch launch of this library.  Only
    // then may the next dispatch carry the programmatic-dependent-launch attribute: it can start
    // (and load its requests and the tables) before its predecessor has finished, which is only
    // sound when the predecessor wrote none of that - dispatch kernels write replies only.
    // Anything else the library enqueues there (copies, memsets, assign_seq, codec, checksum and
    // counter kernels, waits for a new image) clears it; producers the library cannot see are
    // declared by the caller with PTYPE_F_AFTER_PRODUCER.
    std::atomic<bool> after_dispatch{false};
    cudaStream_t copy_stream = nullptr;  // image uploads / broadcast
    cudaStream_t pipe[kPipe] = {};
    Stage stage[kPipe];
    void* image_buf[2] = {nullptr, nullptr};
    size_t image_cap[2] = {0, 0};
    ptype::DeviceImage view[2];
    ptype::LaunchPlan plan[2][kPlanKinds][2];  // [buffer][variant][extras]
    uint64_t epoch[2] = {0, 0};
    int active = -1;
    unsigned long long* d_stats = nullptr;
    uint32_t* d_failed = nullptr;
    size_t failed_cap = 0;
    unsigned long long* d_cksum_partials = nullptr;  // scratch of the checksum kernel
    unsigned long long* d_cksum_out = nullptr;
    unsigned long long* d_size = nullptr;  // broadcast header scratch
    ptype_failed* d_failed2 = nullptr;     // (index, status) list of call_batch_values
    size_t failed2_cap = 0;
    // Balancer counters (rank.cu): one per table slot and image buffer, carried across flips.
    unsigned long long* seq_counters[2] = {nullptr, nullptr};
    size_t seq_cap[2] = {0, 0};             // entries
    bool seq_in_use = false;                // ptype_b200_assign_seq / seq_set has been called
    uint32_t* rank_rows = nullptr;          // (chunks x n_slots) histogram matrix of assign_seq
    size_t rank_rows_cap = 0;               // entries
    uint32_t* rank_ids = nullptr;           // the slot each record of the batch is counted under
    size_t rank_ids_cap = 0;                // entries
    unsigned long long* seq_start = nullptr;
    size_t seq_start_cap = 0;
    unsigned long long* d_tally = nullptr;  // {failed calls, finished CTAs} of a single-launch values batch;
                                            // the kernel leaves both at zero
    unsigned long long* h_tally = nullptr;  // page-locked word the kernel's last CTA writes the count to
    unsigned long long* h_tally_dev = nullptr;
    uint8_t* d_wire_defs = nullptr;        // gob type definitions of rpc.Response and struct{} (wire_encode)
    uint32_t wire_ids[2] = {0, 0}, wire_def_len[2] = {0, 0};
    void* d_codec = nullptr;               // scratch of the gob encoder (lengths, block sums)
    size_t codec_cap = 0;
    // retired[b][i]: recorded on dispatch stream i at the moment image buffer b stopped being the
    // active one - every kernel that reads b was enqueued before it.  Refilling b waits for these
    // only, not for the batches queued since (they read the other buffer), so an upload or
    // broadcast overlaps the dispatches in flight.
    cudaEvent_t retired[2][1 + kPipe] = {};
    bool retired_set[2] = {false, false};
    cudaEvent_t image_ready = nullptr;     // the inactive buffer holds the new image
    void* h_stage[2] = {nullptr, nullptr}; // page-locked copies of the published blobs' device part
    size_t h_stage_cap[2] = {0, 0};
    std::mutex mu;
};

}  // namespace

namespace ptype {
void set_last_error(const char* text) { g_last_error = text ? text : ""; }
}  // namespace ptype

struct ptype_ctx {
    std::vector<DeviceState*> devs;
    std::atomic<uint64_t> check_limit{1ull << 16};
    std::atomic<uint64_t> launches{0};
    void* nccl_comm = nullptr;
    int rank = 0, nranks = 1;
    int comm_spare_sms = 0;  // SMs the dispatch grids leave to the broadcast once updates arrive
    int broadcasts = 0;
    std::mutex image_mu;
    std::vector<unsigned char> host_image;  // last installed blob (host-side lookups)
    bool host_checksum_stale = false;       // update_service patched it since the checksum was taken
};

namespace {

// 128-bit loads and stores need 16-byte aligned record arrays (cudaMalloc, cudaHostAlloc and
// ptype_b200_alloc_pinned all give at least that); 8 for int64 arrays.
bool misaligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) != 0; }

// May the dispatch launch that is about to go onto d->stream overlap its predecessor?
bool take_pdl(DeviceState* d, uint32_t flags) {
    const bool was = d->after_dispatch.exchange(true);
    return was && !(flags & PTYPE_F_AFTER_PRODUCER);
}

DeviceState* dev_of(ptype_ctx* ctx, int dev) {
    if (!ctx || dev < 0 || dev >= static_cast<int>(ctx->devs.size())) return nullptr;
    return ctx->devs[dev];
}

int sync_all_streams(DeviceState* d) {
    CU(cudaStreamSynchronize(d->stream));
    for (int i = 0; i < kPipe; ++i) {
        CU(cudaStreamSynchronize(d->pipe[i]));
        d->stage[i].busy = false;
    }
    return PTYPE_OK;
}

// Snapshot updates do not stop the dispatch streams.  Before the inactive buffer is
// overwritten the copy stream waits (on the device) for everything enqueued so far, since
// launches made two flips ago may still read it; after the upload every dispatch stream
// waits (on the device) for the copy before its next launch.  The host never blocks.
// Before the copy stream overwrites image buffer `buf`: wait for the kernels that may still read it.
int copy_stream_waits_for_readers(DeviceState* d, int buf) {
    if (!d->retired_set[buf]) return PTYPE_OK;  // never been active
    for (int i = 0; i <= kPipe; ++i) CU(cudaStreamWaitEvent(d->copy_stream, d->retired[buf][i], 0));
    return PTYPE_OK;
}
// The flip: dispatches enqueued from here on read buffer `buf`; the buffer they read until now
// is marked retired on every dispatch stream.
// Zeroed counter array of at least n_slots entries for image buffer `buf`.
int ensure_counters(DeviceState* d, int buf, uint32_t n_slots) {
    if (d->seq_cap[buf] >= n_slots) return PTYPE_OK;
    CU(cudaStreamSynchronize(d->stream));
    if (d->seq_counters[buf]) CU(cudaFree(d->seq_counters[buf]));
    d->seq_counters[buf] = nullptr;
    d->seq_cap[buf] = 0;
    CU(cudaMalloc(&d->seq_counters[buf], static_cast<size_t>(n_slots) * sizeof(unsigned long long)));
    CU(cudaMemset(d->seq_counters[buf], 0, static_cast<size_t>(n_slots) * sizeof(unsigned long long)));
    d->seq_cap[buf] = n_slots;
    return PTYPE_OK;
}

int make_active(DeviceState* d, int buf) {
    const int old = d->active;
    d->after_dispatch = false;  // the next dispatch reads tables another stream has just written
    if (old >= 0 && old != buf && d->seq_in_use) {
        // the balancer counters follow their services into the new table, in stream order
        // after every assign_seq enqueued so far (d->stream already waits for the new image)
        int rc = ensure_counters(d, buf, d->view[buf].n_slots);
        if (rc != PTYPE_OK) return rc;
        CU(ptype::launch_carry_counters(d->view[old], d->seq_counters[old], d->view[buf], d->seq_counters[buf],
                                        d->stream));
    }
    if (old >= 0 && old != buf) {
        for (int i = 0; i <= kPipe; ++i) {
            cudaStream_t s = i == 0 ? d->stream : d->pipe[i - 1];
            if (!d->retired[old][i]) CU(cudaEventCreateWithFlags(&d->retired[old][i], cudaEventDisableTiming));
            CU(cudaEventRecord(d->retired[old][i], s));
        }
        d->retired_set[old] = true;
    }
    d->active = buf;
    return PTYPE_OK;
}

int dispatches_wait_for_copy_stream(DeviceState* d) {
    if (!d->image_ready) CU(cudaEventCreateWithFlags(&d->image_ready, cudaEventDisableTiming));
    CU(cudaEventRecord(d->image_ready, d->copy_stream));
    CU(cudaStreamWaitEvent(d->stream, d->image_ready, 0));
    for (int i = 0; i < kPipe; ++i) CU(cudaStreamWaitEvent(d->pipe[i], d->image_ready, 0));
    return PTYPE_OK;
}

// Make the inactive image buffer at least `bytes` big.  Nothing in flight reads it:
// every dispatch launched so far captured the active buffer, and callers sync first.
int reserve_inactive(DeviceState* d, int buf, size_t bytes) {
    if (d->image_cap[buf] >= bytes) return PTYPE_OK;
    int rc0 = sync_all_streams(d);  // growing frees the old buffer: nothing may still read it
    if (rc0 != PTYPE_OK) return rc0;
    if (d->image_buf[buf]) CU(cudaFree(d->image_buf[buf]));
    d->image_buf[buf] = nullptr;
    d->image_cap[buf] = 0;
    size_t cap = (bytes + 4095) / 4096 * 4096;
    CU(cudaMalloc(&d->image_buf[buf], cap));
    d->image_cap[buf] = cap;
    return PTYPE_OK;
}

// Point a DeviceImage at buffer `buf` (which holds at least header.device_bytes of
// the blob), refresh launch plans, and flip.  Mirrors the swap under lock of
// handleNewNodes (cluster/rpc.go:233-236).
int adopt_image(ptype_ctx* ctx, DeviceState* d, int buf, const ptype_image_header& h, const void* host_blob) {
    const unsigned char* base = static_cast<const unsigned char*>(d->image_buf[buf]);
    ptype::DeviceImage& v = d->view[buf];
    v.slots = reinterpret_cast<const ptype_slot*>(base + h.off_slots);
    v.conn_hash = reinterpret_cast<const uint32_t*>(base + h.off_conn_hash);
    v.methods = reinterpret_cast<const ptype_method*>(base + h.off_methods);
    v.mtable = reinterpret_cast<const ptype_method*>(base + h.off_mtable);
    v.table_seed = h.table_seed;
    v.method_mul = h.method_mul;
    v.n_slots = h.n_slots;
    v.max_conn = h.max_connections;
    v.n_methods = h.n_methods;
    // which receivers do the registered services expose? (slot masks, host copy of the blob)
    uint32_t recv_union = 0;
    const ptype_slot* hs = reinterpret_cast<const ptype_slot*>(static_cast<const unsigned char*>(host_blob) + h.off_slots);
    for (uint32_t i = 0; i < h.n_slots; ++i)
        if (hs[i].node_base != PTYPE_SLOT_EMPTY) recv_union |= hs[i].n_and_recv >> 24;
    v.check_only = recv_union == 2u ? 1u : 0u;  // bit 1 = Prime, bit 0 = Calculator (image.cpp)
    v.m_magic = h.max_connections ? ~0ull / h.max_connections + 1ull : 0ull;
    v.check_limit = ctx->check_limit.load();
    for (int g = 0; g < kPlanKinds; ++g)
        for (int x = 0; x < 2; ++x)
            CU(ptype::plan_dispatch(v, kPlanFlags[g], x != 0, &d->plan[buf][g][x]));
    d->epoch[buf] = h.epoch;
    return make_active(d, buf);
}

// Device alias of a page-locked host pointer (UVA maps every cudaHostAlloc / registered
// range), or nullptr for pageable memory.  Lets a kernel stream requests and replies over
// PCIe itself instead of going through chunked copy-engine transfers.
// Grid of a kernel that streams its batch over PCIe: the link, not the SMs, sets the pace, and
// one CTA per SM keeps it fullest (measured: 402 us per 2^20-call batch against 417 us with the
// plan's four, tools/e2e_probe.py).  PTYPE_B200_ZC_CTAS overrides (0 = the plan's grid).
int zero_copy_grid(const ptype::LaunchPlan& plan, size_t n, int sm_count) {
    static const int ctas = [] {
        const char* e = std::getenv("PTYPE_B200_ZC_CTAS");
        return e ? std::atoi(e) : 1;
    }();
    const int full = ptype::grid_for(plan, n, sm_count);
    if (ctas <= 0 || ctas >= plan.ctas_per_sm) return full;
    const int want = sm_count * ctas;
    return want < full ? want : full;
}

void* pinned_alias(const void* host) {
    if (!host) return nullptr;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, host) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

bool single_via_multi() {  // PTYPE_B200_SINGLE_VIA_MULTI=0: single batches keep dispatch_kernel (measurement switch)
    static const bool on = [] {
        const char* v = std::getenv("PTYPE_B200_SINGLE_VIA_MULTI");
        return !(v && v[0] == '0');
    }();
    return on;
}

bool zero_copy_enabled() {
    static const bool on = [] {
        const char* v = std::getenv("PTYPE_B200_ZERO_COPY");
        return !(v && v[0] == '0');
    }();
    return on;
}

int ensure_extras(DeviceState* d, size_t failed_cap) {
    if (!d->d_stats) CU(cudaMalloc(&d->d_stats, ptype::kStatWords * sizeof(unsigned long long)));
    if (failed_cap > d->failed_cap) {
        if (d->d_failed) CU(cudaFree(d->d_failed));
        d->d_failed = nullptr;
        d->failed_cap = 0;
        CU(cudaMalloc(&d->d_failed, failed_cap * sizeof(uint32_t)));
        d->failed_cap = failed_cap;
    }
    return PTYPE_OK;
}

int ensure_stage(DeviceState* d, int s) {
    Stage& st = d->stage[s];
    if (!st.d_reqs) CU(cudaMalloc(&st.d_reqs, chunk_calls() * sizeof(ptype_req)));
    if (!st.d_replies) CU(cudaMalloc(&st.d_replies, chunk_calls() * sizeof(ptype_reply)));
    if (!st.done) CU(cudaEventCreateWithFlags(&st.done, cudaEventDisableTiming));
    return PTYPE_OK;
}

}  // namespace

extern "C" {

const char* ptype_b200_strerror(int err) {
    switch (err) {
        case PTYPE_OK: return "ok";
        case PTYPE_E_INVAL: return "invalid argument";
        case PTYPE_E_NOMEM: return "out of memory";
        case PTYPE_E_CUDA: return "CUDA call failed (see ptype_b200_last_error)";
        case PTYPE_E_NO_IMAGE: return "no registry image installed";
        case PTYPE_E_BAD_IMAGE: return "malformed registry image";
        case PTYPE_E_COLLISION: return "service-key collision could not be resolved";
        case PTYPE_E_NCCL: return "NCCL call failed (see ptype_b200_last_error)";
        case PTYPE_E_LIMIT: return "size beyond a documented maximum";
        case PTYPE_E_NOT_FOUND: return "not found";
        case PTYPE_E_BUSY: return "busy";
        default: return "unknown error";
    }
}

const char* ptype_b200_last_error(void) { return g_last_error.c_str(); }
int ptype_b200_abi_version(void) { return PTYPE_B200_ABI_VERSION; }

int ptype_b200_device_count(int* out) {
    if (!out) return PTYPE_E_INVAL;
    CU(cudaGetDeviceCount(out));
    return PTYPE_OK;
}

void ptype_b200_close(ptype_ctx* ctx) {
    if (!ctx) return;
    if (ctx->nccl_comm) {
        NcclApi* a = nccl_api();
        if (a && a->CommDestroy) a->CommDestroy(ctx->nccl_comm);
    }
    for (DeviceState* d : ctx->devs) {
        if (!d) continue;
        cudaSetDevice(d->device);
        cudaDeviceSynchronize();
        for (int i = 0; i < kPipe; ++i) {
            if (d->stage[i].d_reqs) cudaFree(d->stage[i].d_reqs);
            if (d->stage[i].d_replies) cudaFree(d->stage[i].d_replies);
            if (d->stage[i].done) cudaEventDestroy(d->stage[i].done);
            if (d->pipe[i]) cudaStreamDestroy(d->pipe[i]);
        }
        for (int b = 0; b < 2; ++b)
            if (d->image_buf[b]) cudaFree(d->image_buf[b]);
        if (d->d_stats) cudaFree(d->d_stats);
        if (d->d_failed) cudaFree(d->d_failed);
        if (d->d_cksum_partials) cudaFree(d->d_cksum_partials);
        if (d->d_cksum_out) cudaFree(d->d_cksum_out);
        if (d->d_size) cudaFree(d->d_size);
        if (d->d_codec) cudaFree(d->d_codec);
        if (d->d_wire_defs) cudaFree(d->d_wire_defs);
        if (d->d_failed2) cudaFree(d->d_failed2);
        if (d->d_tally) cudaFree(d->d_tally);
        for (int b = 0; b < 2; ++b)
            if (d->seq_counters[b]) cudaFree(d->seq_counters[b]);
        if (d->rank_rows) cudaFree(d->rank_rows);
        if (d->rank_ids) cudaFree(d->rank_ids);
        if (d->seq_start) cudaFree(d->seq_start);
        if (d->h_tally) cudaFreeHost(d->h_tally);
        for (int i = 0; i <= kPipe; ++i)
            for (int b = 0; b < 2; ++b)
                if (d->retired[b][i]) cudaEventDestroy(d->retired[b][i]);
        if (d->image_ready) cudaEventDestroy(d->image_ready);
        for (int b = 0; b < 2; ++b)
            if (d->h_stage[b]) cudaFreeHost(d->h_stage[b]);
        if (d->stream) cudaStreamDestroy(d->stream);
        if (d->copy_stream) cudaStreamDestroy(d->copy_stream);
        delete d;
    }
    delete ctx;
}

int ptype_b200_open(const int* devices, int ndev, ptype_ctx** out) {
    if (!out || ndev < 1 || !devices) return PTYPE_E_INVAL;
    int count = 0;
    CU(cudaGetDeviceCount(&count));
    ptype_ctx* ctx = new (std::nothrow) ptype_ctx();
    if (!ctx) return PTYPE_E_NOMEM;
    for (int i = 0; i < ndev; ++i) {
        if (devices[i] < 0 || devices[i] >= count) {
            ptype_b200_close(ctx);
            return PTYPE_E_INVAL;
        }
        DeviceState* d = new (std::nothrow) DeviceState();
        if (!d) {
            ptype_b200_close(ctx);
            return PTYPE_E_NOMEM;
        }
        ctx->devs.push_back(d);
        d->device = devices[i];
        cudaError_t e = cudaSetDevice(d->device);
        if (e == cudaSuccess)
            e = cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, d->device);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking);
        // uploads and the snapshot broadcast go first wherever they compete with dispatch CTAs
        int prio_least = 0, prio_greatest = 0;
        if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
        if (e == cudaSuccess)
            e = cudaStreamCreateWithPriority(&d->copy_stream, cudaStreamNonBlocking, prio_greatest);
        d->dispatch_sms = d->sm_count;
        for (int s = 0; s < kPipe && e == cudaSuccess; ++s)
            e = cudaStreamCreateWithFlags(&d->pipe[s], cudaStreamNonBlocking);
        if (e != cudaSuccess) {
            ptype_b200_close(ctx);
            return cuda_fail(e, "ptype_b200_open");
        }
    }
    *out = ctx;
    return PTYPE_OK;
}

void* ptype_b200_alloc_pinned(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void ptype_b200_free_pinned(void* p) {
    if (p) cudaFreeHost(p);
}

int ptype_b200_device_alloc(ptype_ctx* ctx, int dev, size_t bytes, void** out) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !out) return PTYPE_E_INVAL;
    CU(cudaSetDevice(d->device));
    CU(cudaMalloc(out, bytes ? bytes : 1));
    return PTYPE_OK;
}
int ptype_b200_device_free(ptype_ctx* ctx, int dev, void* p) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d) return PTYPE_E_INVAL;
    CU(cudaSetDevice(d->device));
    CU(cudaFree(p));
    return PTYPE_OK;
}
int ptype_b200_memcpy_h2d(ptype_ctx* ctx, int dev, void* dst, const void* src, size_t bytes) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d) return PTYPE_E_INVAL;
    CU(cudaSetDevice(d->device));
    d->after_dispatch = false;
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}
int ptype_b200_memcpy_d2h(ptype_ctx* ctx, int dev, void* dst, const void* src, size_t bytes) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d) return PTYPE_E_INVAL;
    CU(cudaSetDevice(d->device));
    d->after_dispatch = false;
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}
int ptype_b200_stream(ptype_ctx* ctx, int dev, void** cuda_stream) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !cuda_stream) return PTYPE_E_INVAL;
    *cuda_stream = d->stream;
    d->after_dispatch = false;
    return PTYPE_OK;
}

int ptype_b200_set_check_limit(ptype_ctx* ctx, uint64_t max_iterations) {
    if (!ctx) return PTYPE_E_INVAL;
    ctx->check_limit.store(max_iterations);
    for (DeviceState* d : ctx->devs) {
        std::lock_guard<std::mutex> lk(d->mu);
        d->view[0].check_limit = max_iterations;
        d->view[1].check_limit = max_iterations;
    }
    return PTYPE_OK;
}

int ptype_b200_publish_image(ptype_ctx* ctx, const void* image, size_t bytes) {
    if (!ctx) return PTYPE_E_INVAL;
    int rc = ptype_b200_image_validate(image, bytes);
    if (rc != PTYPE_OK) return rc;
    const ptype_image_header& h = *static_cast<const ptype_image_header*>(image);
    for (DeviceState* d : ctx->devs) {
        std::lock_guard<std::mutex> lk(d->mu);
        CU(cudaSetDevice(d->device));
        const int buf = d->active == 0 ? 1 : 0;
        rc = reserve_inactive(d, buf, h.device_bytes);
        if (rc != PTYPE_OK) return rc;
        // Stage the blob in page-locked memory so the upload is a true asynchronous DMA.  The
        // staging slot was last used two publishes ago; its copy has long been ordered before
        // the previous publish's, which the copy stream has to have finished first.
        if (d->h_stage_cap[buf] < h.device_bytes) {
            CU(cudaStreamSynchronize(d->copy_stream));
            if (d->h_stage[buf]) CU(cudaFreeHost(d->h_stage[buf]));
            d->h_stage[buf] = nullptr;
            d->h_stage_cap[buf] = 0;
            CU(cudaHostAlloc(&d->h_stage[buf], (h.device_bytes + 4095) / 4096 * 4096, cudaHostAllocDefault));
            d->h_stage_cap[buf] = (h.device_bytes + 4095) / 4096 * 4096;
        } else {
            CU(cudaStreamSynchronize(d->copy_stream));
        }
        std::memcpy(d->h_stage[buf], image, h.device_bytes);
        rc = copy_stream_waits_for_readers(d, buf);
        if (rc != PTYPE_OK) return rc;
        CU(cudaMemcpyAsync(d->image_buf[buf], d->h_stage[buf], h.device_bytes, cudaMemcpyHostToDevice,
                           d->copy_stream));
        rc = dispatches_wait_for_copy_stream(d);
        if (rc != PTYPE_OK) return rc;
        rc = adopt_image(ctx, d, buf, h, image);  // the flip: later dispatches use the new image
        if (rc != PTYPE_OK) return rc;
    }
    std::lock_guard<std::mutex> lk(ctx->image_mu);
    ctx->host_image.assign(static_cast<const unsigned char*>(image),
                           static_cast<const unsigned char*>(image) + bytes);
    ctx->host_checksum_stale = false;
    return PTYPE_OK;
}

int ptype_b200_update_service(ptype_ctx* ctx, const char* service, int delta) {
    if (!ctx || !service) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> ilk(ctx->image_mu);
    if (ctx->host_image.empty()) return PTYPE_E_NO_IMAGE;
    ptype_image_header* hh = reinterpret_cast<ptype_image_header*>(ctx->host_image.data());
    const uint64_t key = ptype_b200_service_key(service, hh->key_salt);
    // Patch the host copy first: it tells whether the service exists and keeps host-side readers
    // of the device part in step.  Same rule as the kernel (update.cu).
    ptype_slot* hs = reinterpret_cast<ptype_slot*>(ctx->host_image.data() + hh->off_slots);
    uint32_t at = PTYPE_SLOT_EMPTY;
    for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
        const uint32_t c = ptype::slot_of(static_cast<uint32_t>(key), static_cast<uint32_t>(key >> 32),
                                          hh->table_seed, w, hh->n_slots);
        if (hs[c].node_base != PTYPE_SLOT_EMPTY && hs[c].key == key) at = c;
    }
    if (at == PTYPE_SLOT_EMPTY) return PTYPE_E_NOT_FOUND;
    const uint32_t n_old = hs[at].n_and_recv & 0xFFFFFFu;
    const long long n_new = static_cast<long long>(n_old) + delta;
    if (n_new < 0 || n_new > 0xFFFFFF) return PTYPE_E_LIMIT;
    for (DeviceState* d : ctx->devs) {
        std::lock_guard<std::mutex> lk(d->mu);
        if (d->active < 0) return PTYPE_E_NO_IMAGE;
        CU(cudaSetDevice(d->device));
        const int src = d->active, buf = src == 0 ? 1 : 0;
        int rc = reserve_inactive(d, buf, hh->device_bytes);
        if (rc != PTYPE_OK) return rc;
        if (!d->d_size) CU(cudaMalloc(&d->d_size, 16));
        rc = copy_stream_waits_for_readers(d, buf);
        if (rc != PTYPE_OK) return rc;
        CU(cudaMemcpyAsync(d->image_buf[buf], d->image_buf[src], hh->device_bytes, cudaMemcpyDeviceToDevice,
                           d->copy_stream));
        CU(ptype::launch_patch_service(
            reinterpret_cast<ptype_slot*>(static_cast<unsigned char*>(d->image_buf[buf]) + hh->off_slots),
            reinterpret_cast<const uint32_t*>(static_cast<unsigned char*>(d->image_buf[buf]) + hh->off_order),
            hh->n_slots, hh->table_seed, key, delta, reinterpret_cast<uint32_t*>(d->d_size), d->copy_stream));
        ctx->launches.fetch_add(1);
        rc = dispatches_wait_for_copy_stream(d);
        if (rc != PTYPE_OK) return rc;
        // same tables, same plans: only the pointers and the epoch move
        const unsigned char* base = static_cast<const unsigned char*>(d->image_buf[buf]);
        ptype::DeviceImage v = d->view[src];
        v.slots = reinterpret_cast<const ptype_slot*>(base + hh->off_slots);
        v.conn_hash = reinterpret_cast<const uint32_t*>(base + hh->off_conn_hash);
        v.methods = reinterpret_cast<const ptype_method*>(base + hh->off_methods);
        v.mtable = reinterpret_cast<const ptype_method*>(base + hh->off_mtable);
        d->view[buf] = v;
        for (int g = 0; g < kPlanKinds; ++g)
            for (int x = 0; x < 2; ++x) d->plan[buf][g][x] = d->plan[src][g][x];
        d->epoch[buf] = d->epoch[src] + 1;
        rc = make_active(d, buf);
        if (rc != PTYPE_OK) return rc;
    }
    const uint32_t* order = reinterpret_cast<const uint32_t*>(ctx->host_image.data() + hh->off_order);
    for (uint32_t i = 0; i < hh->n_slots; ++i) {
        if (hs[i].node_base == PTYPE_SLOT_EMPTY) continue;
        if (i == at)
            hs[i].n_and_recv = (hs[i].n_and_recv & 0xFF000000u) | static_cast<uint32_t>(n_new);
        else if (order[i] > order[at])
            hs[i].node_base = static_cast<uint32_t>(static_cast<long long>(hs[i].node_base) + delta);
    }
    // the blob no longer matches its node/string sections: keep only what the kernels read
    hh->epoch += 1;
    hh->n_nodes = static_cast<uint32_t>(static_cast<long long>(hh->n_nodes) + delta);
    hh->host_sections = 0;
    ctx->host_checksum_stale = true;  // recomputed when somebody asks for the blob
    return PTYPE_OK;
}

int ptype_b200_epoch(ptype_ctx* ctx, int dev, uint64_t* epoch) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !epoch) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    *epoch = d->epoch[d->active];
    return PTYPE_OK;
}

// Copy of the blob last installed in this context (on non-root ranks: the bytes
// that arrived over NCCL).  *bytes in: capacity, out: size needed.
int ptype_b200_current_image(ptype_ctx* ctx, void* out, size_t* bytes) {
    if (!ctx || !bytes) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(ctx->image_mu);
    if (ctx->host_image.empty()) return PTYPE_E_NO_IMAGE;
    if (ctx->host_checksum_stale) {
        ptype_image_header* hh = reinterpret_cast<ptype_image_header*>(ctx->host_image.data());
        hh->body_checksum = ptype::image_body_hash(ctx->host_image.data() + sizeof(ptype_image_header),
                                                   ctx->host_image.size() - sizeof(ptype_image_header));
        ctx->host_checksum_stale = false;
    }
    const size_t need = ctx->host_image.size();
    if (out && *bytes >= need) std::memcpy(out, ctx->host_image.data(), need);
    const bool fits = out && *bytes >= need;
    *bytes = need;
    return fits ? PTYPE_OK : PTYPE_E_LIMIT;
}

int ptype_b200_nccl_unique_id(void* id128) {
    NcclApi* a = nccl_api();
    if (!a) {
        g_last_error = "libnccl.so.2 not loadable";
        return PTYPE_E_NCCL;
    }
    int r = a->GetUniqueId(id128);
    return r == 0 ? PTYPE_OK : nccl_fail(r, "ncclGetUniqueId");
}

int ptype_b200_comm_init(ptype_ctx* ctx, const void* id128, int rank, int nranks) {
    if (!ctx || !id128 || ctx->devs.size() != 1 || rank < 0 || rank >= nranks) return PTYPE_E_INVAL;
    NcclApi* a = nccl_api();
    if (!a) {
        g_last_error = "libnccl.so.2 not loadable";
        return PTYPE_E_NCCL;
    }
    CU(cudaSetDevice(ctx->devs[0]->device));
    UniqueId128 id;
    std::memcpy(&id, id128, sizeof id);
    void* comm = nullptr;
    int r = a->CommInitRank(&comm, nranks, id, rank);
    if (r != 0) return nccl_fail(r, "ncclCommInitRank");
    ctx->nccl_comm = comm;
    ctx->rank = rank;
    ctx->nranks = nranks;
    const char* e = std::getenv("PTYPE_B200_COMM_SMS");
    ctx->comm_spare_sms = e ? std::atoi(e) : 1;
    if (ctx->comm_spare_sms < 0) ctx->comm_spare_sms = 0;
    return PTYPE_OK;
}

// Registry-snapshot broadcast: the only traffic that crosses NVLink on this path.
int ptype_b200_broadcast_image(ptype_ctx* ctx, const void* image, size_t bytes, int root) {
    if (!ctx || !ctx->nccl_comm || ctx->devs.size() != 1) return PTYPE_E_INVAL;
    NcclApi* a = nccl_api();
    DeviceState* d = ctx->devs[0];
    const bool is_root = ctx->rank == root;
    if (is_root) {
        int rc = ptype_b200_image_validate(image, bytes);
        if (rc != PTYPE_OK) return rc;
    }
    std::lock_guard<std::mutex> lk2(ctx->image_mu);  // lock order: image_mu, then a device's mu
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    // The dispatch kernels are persistent and fill every SM, so a collective's kernel enqueued
    // beside them does not start until the running launch (up to 96 batches) has drained, and
    // this call blocks the host until then.  A context that receives live updates - this is its
    // second broadcast - sizes its dispatch grids for PTYPE_B200_COMM_SMS fewer SMs (default 1)
    // from here on; that is where ncclBroadcast runs while the batches in flight carry on
    // (measured, 2 GPUs, an update every 64 batches: 135.7 -> 162.7 G calls/s).
    if (++ctx->broadcasts == 2 && d->sm_count - ctx->comm_spare_sms >= 1)
        d->dispatch_sms = d->sm_count - ctx->comm_spare_sms;
    if (!d->d_size) CU(cudaMalloc(&d->d_size, 16));
    unsigned long long total = is_root ? bytes : 0;
    if (is_root)
        CU(cudaMemcpyAsync(d->d_size, &total, sizeof total, cudaMemcpyHostToDevice, d->copy_stream));
    int r = a->Broadcast(d->d_size, d->d_size, sizeof total, kNcclUint8, root, ctx->nccl_comm,
                         d->copy_stream);
    if (r != 0) return nccl_fail(r, "ncclBroadcast(size)");
    CU(cudaMemcpyAsync(&total, d->d_size, sizeof total, cudaMemcpyDeviceToHost, d->copy_stream));
    CU(cudaStreamSynchronize(d->copy_stream));
    if (total < sizeof(ptype_image_header) || total > (1ull << 32)) return PTYPE_E_BAD_IMAGE;
    const int buf = d->active == 0 ? 1 : 0;
    int rc = reserve_inactive(d, buf, total);
    if (rc != PTYPE_OK) return rc;
    rc = copy_stream_waits_for_readers(d, buf);
    if (rc != PTYPE_OK) return rc;
    if (is_root)
        CU(cudaMemcpyAsync(d->image_buf[buf], image, total, cudaMemcpyHostToDevice, d->copy_stream));
    r = a->Broadcast(d->image_buf[buf], d->image_buf[buf], total, kNcclUint8, root, ctx->nccl_comm,
                     d->copy_stream);
    if (r != 0) return nccl_fail(r, "ncclBroadcast(image)");
    std::vector<unsigned char> host(total);
    if (is_root) {
        std::memcpy(host.data(), image, total);
    } else {
        CU(cudaMemcpyAsync(host.data(), d->image_buf[buf], total, cudaMemcpyDeviceToHost,
                           d->copy_stream));
    }
    CU(cudaStreamSynchronize(d->copy_stream));
    rc = ptype_b200_image_validate(host.data(), host.size());
    if (rc != PTYPE_OK) return rc;
    rc = dispatches_wait_for_copy_stream(d);
    if (rc != PTYPE_OK) return rc;
    rc = adopt_image(ctx, d, buf, *reinterpret_cast<const ptype_image_header*>(host.data()), host.data());
    if (rc != PTYPE_OK) return rc;
    ctx->host_image.swap(host);
    ctx->host_checksum_stale = false;
    return PTYPE_OK;
}

int ptype_b200_sync(ptype_ctx* ctx, int dev) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    return sync_all_streams(d);
}

int ptype_b200_dispatch_ex(ptype_ctx* ctx, int dev, const ptype_req* reqs, size_t n,
                           ptype_reply* replies, uint32_t* failed_idx, size_t failed_cap,
                           ptype_batch_stats* stats, uint32_t flags) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || (n && (!reqs || !replies))) return PTYPE_E_INVAL;
    if (misaligned(reqs, 16) || misaligned(replies, 16) || misaligned(failed_idx, 4) || misaligned(stats, 8))
        return PTYPE_E_INVAL;
    if (n > 0xFFFFFFFFull || failed_cap > 0xFFFFFFFFull) return PTYPE_E_LIMIT;
    if (failed_idx && !stats) return PTYPE_E_INVAL;  // the list's length lives in stats
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    CU(cudaSetDevice(d->device));
    const int buf = d->active;
    const ptype::DeviceImage img = d->view[buf];
    const bool extras = stats != nullptr;
    const ptype::LaunchPlan& plan = d->plan[buf][plan_kind(flags)][extras ? 1 : 0];

    if (flags & PTYPE_F_DEVICE_PTRS) {
        // Batch resident in HBM: one kernel launch, nothing else.
        unsigned long long* d_stats = nullptr;
        if (extras) {
            d_stats = reinterpret_cast<unsigned long long*>(stats);
            CU(cudaMemsetAsync(d_stats, 0, sizeof(ptype_batch_stats), d->stream));
            d->after_dispatch = false;
        }
        if (n) {
            if (!extras && !plan.bulk && single_via_multi()) {
                // one batch is a batch list of one: the same (software-pipelined) loop as dispatch_batches
                const ptype_batch one = {reqs, replies, n};
                CU(ptype::launch_dispatch_multi(img, plan, d->dispatch_sms, &one, 1, d->stream, take_pdl(d, flags)));
            } else {
                CU(ptype::launch_dispatch(img, plan, ptype::grid_for(plan, n, d->dispatch_sms), reqs, n,
                                          replies, failed_idx, static_cast<uint32_t>(failed_cap), 0u,
                                          d_stats, d->stream, take_pdl(d, flags)));
            }
            ctx->launches.fetch_add(1);
        }
        if (!(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
        return PTYPE_OK;
    }

    // Page-locked host buffers: the kernel reads the records and writes the replies across
    // PCIe itself (both directions overlap, no staging, one launch).
    if (zero_copy_enabled() && !extras && n) {
        const ptype_req* q = static_cast<const ptype_req*>(pinned_alias(reqs));
        ptype_reply* r = static_cast<ptype_reply*>(pinned_alias(replies));
        if (q && r) {
            CU(ptype::launch_dispatch(img, plan, zero_copy_grid(plan, n, d->dispatch_sms), q, n, r, nullptr,
                                      0u, 0u, nullptr, d->stream, take_pdl(d, flags)));
            ctx->launches.fetch_add(1);
            if (!(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
            return PTYPE_OK;
        }
    }
    // Host buffers: chunked H2D -> kernel -> D2H pipeline over kPipe streams.
    int rc;
    if (extras) {
        rc = ensure_extras(d, failed_idx ? failed_cap : 0);
        if (rc != PTYPE_OK) return rc;
        CU(cudaMemsetAsync(d->d_stats, 0, ptype::kStatWords * sizeof(unsigned long long), d->pipe[0]));
        CU(cudaStreamSynchronize(d->pipe[0]));
    }
    size_t done = 0;
    int s = 0;
    while (done < n) {
        const size_t cnt = n - done < chunk_calls() ? n - done : chunk_calls();
        rc = ensure_stage(d, s);
        if (rc != PTYPE_OK) return rc;
        Stage& st = d->stage[s];
        if (st.busy) {
            CU(cudaEventSynchronize(st.done));
            st.busy = false;
        }
        CU(cudaMemcpyAsync(st.d_reqs, reqs + done, cnt * sizeof(ptype_req), cudaMemcpyHostToDevice,
                           d->pipe[s]));
        CU(ptype::launch_dispatch(img, plan, ptype::grid_for(plan, cnt, d->dispatch_sms), st.d_reqs, cnt,
                                  st.d_replies, failed_idx ? d->d_failed : nullptr,
                                  static_cast<uint32_t>(failed_cap), static_cast<uint32_t>(done),
                                  extras ? d->d_stats : nullptr, d->pipe[s], false));
        ctx->launches.fetch_add(1);
        CU(cudaMemcpyAsync(replies + done, st.d_replies, cnt * sizeof(ptype_reply),
                           cudaMemcpyDeviceToHost, d->pipe[s]));
        CU(cudaEventRecord(st.done, d->pipe[s]));
        st.busy = true;
        done += cnt;
        s = (s + 1) % kPipe;
    }
    if (extras || !(flags & PTYPE_F_ASYNC)) {
        for (int i = 0; i < kPipe; ++i) {
            if (d->stage[i].busy) {
                CU(cudaEventSynchronize(d->stage[i].done));
                d->stage[i].busy = false;
            }
        }
    }
    if (extras) {
        unsigned long long words[ptype::kStatWords];
        CU(cudaMemcpyAsync(words, d->d_stats, sizeof words, cudaMemcpyDeviceToHost, d->pipe[0]));
        CU(cudaStreamSynchronize(d->pipe[0]));
        stats->calls = words[ptype::kStatCalls];
        for (int i = 0; i < PTYPE_ST__COUNT; ++i) stats->by_status[i] = words[ptype::kStatByStatus + i];
        stats->failed_count = words[ptype::kStatFailed];
        if (failed_idx && failed_cap) {
            size_t take = stats->failed_count < failed_cap ? stats->failed_count : failed_cap;
            if (take) {
                CU(cudaMemcpyAsync(failed_idx, d->d_failed, take * sizeof(uint32_t),
                                   cudaMemcpyDeviceToHost, d->pipe[0]));
                CU(cudaStreamSynchronize(d->pipe[0]));
            }
        }
    }
    return PTYPE_OK;
}

int ptype_b200_dispatch_batches(ptype_ctx* ctx, int dev, const ptype_batch* batches, size_t count,
                                uint32_t flags) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || (count && !batches)) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    CU(cudaSetDevice(d->device));
    const int buf = d->active;
    const ptype::DeviceImage img = d->view[buf];
    const ptype::LaunchPlan& plan = d->plan[buf][plan_kind(flags)][0];
    for (size_t b = 0; b < count; ++b) {
        if (batches[b].n && (!batches[b].reqs || !batches[b].replies)) return PTYPE_E_INVAL;
        if (misaligned(batches[b].reqs, 16) || misaligned(batches[b].replies, 16)) return PTYPE_E_INVAL;
        if (batches[b].n > 0xFFFFFFFFull) return PTYPE_E_LIMIT;
    }
    if (!(flags & PTYPE_F_ONE_LAUNCH_PER_BATCH)) {
        // One launch per group of batches: the persistent grid walks their tiles as one
        // stream.  Groups are cut so that a launch stays below 2^32 tiles.
        const size_t group = static_cast<size_t>(ptype::max_batches_per_launch());
        size_t b = 0;
        while (b < count) {
            size_t e = b;
            uint64_t tiles = 0;
            while (e < count && e - b < group && tiles + (batches[e].n + 31) / 32 < 0xFFFFFFFFull) {
                tiles += (batches[e].n + 31) / 32;
                ++e;
            }
            if (e == b) return PTYPE_E_LIMIT;
            if (tiles) {
                // a list that finds the stream idle is on its own: its end may run as a second launch
                // (cudaStreamQuery reports cudaErrorNotReady for a busy stream; that is not a failure)
                bool lone = false;
                if (e - b >= 2) {
                    lone = cudaStreamQuery(d->stream) == cudaSuccess;
                    if (!lone) cudaGetLastError();  // leave no "not ready" behind for a later error check
                }
                int kernels = 0;
                CU(ptype::launch_dispatch_multi(img, plan, d->dispatch_sms, batches + b, static_cast<int>(e - b),
                                                d->stream, take_pdl(d, flags), nullptr, &kernels, lone));
                ctx->launches.fetch_add(static_cast<uint64_t>(kernels));
            }
            b = e;
        }
    } else {
        for (size_t b = 0; b < count; ++b) {
            const ptype_batch& bt = batches[b];
            if (bt.n == 0) continue;
            CU(ptype::launch_dispatch(img, plan, ptype::grid_for(plan, bt.n, d->dispatch_sms), bt.reqs, bt.n,
                                      bt.replies, nullptr, 0u, 0u, nullptr, d->stream, take_pdl(d, flags)));
            ctx->launches.fetch_add(1);
        }
    }
    if (!(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}

namespace {

// Shared body of ptype_b200_call_batch (16-byte reply records) and
// ptype_b200_call_batch_values (8-byte values + failed list).
int call_batch_impl(ptype_ctx* ctx, int dev, uint64_t svc_key, uint64_t seq_base,
                    const ptype_method_ref* methods, int n_methods, const uint8_t* method_sel,
                    const int64_t* args, int args_per_call, size_t n, ptype_reply* replies,
                    int64_t* values, ptype_failed* failed, size_t failed_cap, uint64_t* failed_count,
                    uint32_t flags) {
    DeviceState* d = dev_of(ctx, dev);
    const bool values_mode = values != nullptr;
    if (!d || !methods || n_methods < 1 || n_methods > PTYPE_BATCH_METHODS ||
        (args_per_call != 2 && args_per_call != 3) || (n && (!args || (!replies && !values))) ||
        (failed_cap && !failed))
        return PTYPE_E_INVAL;
    if (misaligned(args, args_per_call == 2 ? 16 : 8) || misaligned(replies, 16) || misaligned(values, 8))
        return PTYPE_E_INVAL;
    if (n > 0xFFFFFFFFull || failed_cap > 0xFFFFFFFFull) return PTYPE_E_LIMIT;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    CU(cudaSetDevice(d->device));
    const int buf = d->active;
    const ptype::DeviceImage img = d->view[buf];
    // the batch kernel loads with LDG: use that plan's block shape and table placement
    const ptype::LaunchPlan& plan =
        d->plan[buf][plan_kind((flags & PTYPE_F_GLOBAL_TABLE) | PTYPE_F_LDG_LOADS)][0];
    uint32_t m[PTYPE_BATCH_METHODS], r[PTYPE_BATCH_METHODS];
    for (int i = 0; i < n_methods; ++i) {
        m[i] = methods[i].method;
        r[i] = methods[i].recv;
    }
    ptype::CallBatchOut out;
    // Where the batch goes out as ONE launch (device pointers, or page-locked host buffers the
    // kernel streams itself) the kernel reports the failure count through h_tally; the chunked
    // pipeline below launches several kernels at once and keeps the memset / copy-back protocol.
    bool single_launch = (flags & PTYPE_F_DEVICE_PTRS) != 0;
    const int64_t* za = nullptr;
    ptype_reply* zr = nullptr;
    int64_t* zv = nullptr;
    const uint8_t* zs = nullptr;
    bool zero_copy = false;
    if (!single_launch && zero_copy_enabled() && n) {
        za = static_cast<const int64_t*>(pinned_alias(args));
        zr = replies ? static_cast<ptype_reply*>(pinned_alias(replies)) : nullptr;
        zv = values ? static_cast<int64_t*>(pinned_alias(values)) : nullptr;
        zs = method_sel ? static_cast<const uint8_t*>(pinned_alias(method_sel)) : nullptr;
        zero_copy = za && (values_mode ? zv != nullptr : zr != nullptr) && (!method_sel || zs);
        single_launch = zero_copy;
    }
    if (values_mode) {
        // the failed list and its counter live in device memory; copied out at the end
        if (failed_cap > d->failed2_cap) {
            CU(cudaStreamSynchronize(d->stream));
            if (d->d_failed2) CU(cudaFree(d->d_failed2));
            d->d_failed2 = nullptr;
            d->failed2_cap = 0;
            CU(cudaMalloc(&d->d_failed2, failed_cap * sizeof(ptype_failed)));
            d->failed2_cap = failed_cap;
        }
        out.failed = d->d_failed2;
        out.failed_cap = static_cast<uint32_t>(failed_cap);
        if (single_launch) {
            if (!d->d_tally) {
                CU(cudaMalloc(&d->d_tally, 2 * sizeof(unsigned long long)));
                CU(cudaMemset(d->d_tally, 0, 2 * sizeof(unsigned long long)));
            }
            if (!d->h_tally) {
                CU(cudaHostAlloc(reinterpret_cast<void**>(&d->h_tally), sizeof(unsigned long long),
                                 cudaHostAllocMapped | cudaHostAllocPortable));
                CU(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d->h_tally_dev), d->h_tally, 0));
            }
            *d->h_tally = 0;
            out.failed_count = d->d_tally;
            out.done_ctas = d->d_tally + 1;
            out.host_count = d->h_tally_dev;
        } else {
            if (!d->d_stats) CU(cudaMalloc(&d->d_stats, ptype::kStatWords * sizeof(unsigned long long)));
            CU(cudaMemsetAsync(d->d_stats, 0, sizeof(unsigned long long), d->stream));
            CU(cudaStreamSynchronize(d->stream));
            d->after_dispatch = false;
            out.failed_count = d->d_stats;
        }
    }
    auto finish_values = [&]() -> int {
        if (!values_mode) return PTYPE_OK;
        unsigned long long count = 0;
        if (single_launch)
            count = n ? *static_cast<volatile unsigned long long*>(d->h_tally) : 0;
        else
            CU(cudaMemcpy(&count, d->d_stats, sizeof count, cudaMemcpyDeviceToHost));
        if (failed_count) *failed_count = count;
        const size_t take = count < failed_cap ? static_cast<size_t>(count) : failed_cap;
        if (take) CU(cudaMemcpy(failed, d->d_failed2, take * sizeof(ptype_failed), cudaMemcpyDeviceToHost));
        return PTYPE_OK;
    };
    const size_t arg_bytes = static_cast<size_t>(args_per_call) * sizeof(int64_t);
    if (flags & PTYPE_F_DEVICE_PTRS) {
        if (n) {
            out.values = values;
            CU(ptype::launch_call_batch(img, plan, ptype::grid_for(plan, n, d->dispatch_sms), svc_key, seq_base,
                                        m, r, n_methods, args, args_per_call, method_sel, n, replies, out,
                                        d->stream, take_pdl(d, flags)));
            ctx->launches.fetch_add(1);
        }
        if (values_mode || !(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
        return finish_values();
    }
    if (zero_copy) {
        out.values = zv;
        CU(ptype::launch_call_batch(img, plan, zero_copy_grid(plan, n, d->dispatch_sms), svc_key, seq_base,
                                    m, r, n_methods, za, args_per_call, zs, n, zr, out, d->stream,
                                    take_pdl(d, flags)));
        ctx->launches.fetch_add(1);
        if (values_mode || !(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
        return finish_values();
    }
    // Pageable host buffers: the staging slot of each pipeline stream holds a chunk's arguments
    // (<= 24 B/call) and, behind them, its method selectors (1 B/call); the reply slot holds
    // the chunk's records or values.
    size_t done = 0;
    int s = 0;
    while (done < n) {
        const size_t cnt = n - done < chunk_calls() ? n - done : chunk_calls();
        int rc = ensure_stage(d, s);
        if (rc != PTYPE_OK) return rc;
        Stage& st = d->stage[s];
        if (st.busy) {
            CU(cudaEventSynchronize(st.done));
            st.busy = false;
        }
        unsigned char* base = reinterpret_cast<unsigned char*>(st.d_reqs);
        int64_t* d_args = reinterpret_cast<int64_t*>(base);
        uint8_t* d_sel = method_sel ? base + chunk_calls() * 24 : nullptr;
        CU(cudaMemcpyAsync(d_args, reinterpret_cast<const unsigned char*>(args) + done * arg_bytes,
                           cnt * arg_bytes, cudaMemcpyHostToDevice, d->pipe[s]));
        if (method_sel)
            CU(cudaMemcpyAsync(d_sel, method_sel + done, cnt, cudaMemcpyHostToDevice, d->pipe[s]));
        out.values = values_mode ? reinterpret_cast<int64_t*>(st.d_replies) : nullptr;
        out.index_base = static_cast<uint32_t>(done);
        CU(ptype::launch_call_batch(img, plan, ptype::grid_for(plan, cnt, d->dispatch_sms), svc_key,
                                    seq_base + done, m, r, n_methods, d_args, args_per_call, d_sel, cnt,
                                    st.d_replies, out, d->pipe[s], false));
        ctx->launches.fetch_add(1);
        if (values_mode)
            CU(cudaMemcpyAsync(values + done, st.d_replies, cnt * sizeof(int64_t), cudaMemcpyDeviceToHost,
                               d->pipe[s]));
        else
            CU(cudaMemcpyAsync(replies + done, st.d_replies, cnt * sizeof(ptype_reply),
                               cudaMemcpyDeviceToHost, d->pipe[s]));
        CU(cudaEventRecord(st.done, d->pipe[s]));
        st.busy = true;
        done += cnt;
        s = (s + 1) % kPipe;
    }
    if (values_mode || !(flags & PTYPE_F_ASYNC)) {
        for (int i = 0; i < kPipe; ++i) {
            if (d->stage[i].busy) {
                CU(cudaEventSynchronize(d->stage[i].done));
                d->stage[i].busy = false;
            }
        }
    }
    return finish_values();
}

}  // namespace

int ptype_b200_call_batch(ptype_ctx* ctx, int dev, uint64_t svc_key, uint64_t seq_base,
                          const ptype_method_ref* methods, int n_methods, const uint8_t* method_sel,
                          const int64_t* args, int args_per_call, size_t n, ptype_reply* replies,
                          uint32_t flags) {
    if (n && !replies) return PTYPE_E_INVAL;
    return call_batch_impl(ctx, dev, svc_key, seq_base, methods, n_methods, method_sel, args, args_per_call, n,
                           replies, nullptr, nullptr, 0, nullptr, flags);
}

int ptype_b200_call_batch_values(ptype_ctx* ctx, int dev, uint64_t svc_key, uint64_t seq_base,
                                 const ptype_method_ref* methods, int n_methods, const uint8_t* method_sel,
                                 const int64_t* args, int args_per_call, size_t n, int64_t* values,
                                 ptype_failed* failed, size_t failed_cap, uint64_t* failed_count,
                                 uint32_t flags) {
    if (!values) return PTYPE_E_INVAL;
    if (failed_count) *failed_count = 0;
    if (n == 0) return PTYPE_OK;
    return call_batch_impl(ctx, dev, svc_key, seq_base, methods, n_methods, method_sel, args, args_per_call, n,
                           nullptr, values, failed, failed_cap, failed_count, flags);
}

int ptype_b200_dispatch(ptype_ctx* ctx, int dev, const ptype_req* reqs, size_t n,
                        ptype_reply* replies, uint32_t flags) {
    return ptype_b200_dispatch_ex(ctx, dev, reqs, n, replies, nullptr, 0, nullptr, flags);
}

int ptype_b200_reply_checksum(ptype_ctx* ctx, int dev, const ptype_reply* replies, size_t n,
                              uint64_t index_base, uint64_t* checksum, uint32_t flags) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !checksum || (n && !replies)) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    if (!d->d_cksum_partials)
        CU(cudaMalloc(&d->d_cksum_partials, ptype::kChecksumBlocks * sizeof(unsigned long long)));
    if (!d->d_cksum_out) CU(cudaMalloc(&d->d_cksum_out, sizeof(unsigned long long)));
    const ptype_reply* d_replies = replies;
    void* tmp = nullptr;
    d->after_dispatch = false;
    if (!(flags & PTYPE_F_DEVICE_PTRS) && n) {
        CU(cudaMalloc(&tmp, n * sizeof(ptype_reply)));
        cudaError_t e = cudaMemcpyAsync(tmp, replies, n * sizeof(ptype_reply), cudaMemcpyHostToDevice,
                                        d->stream);
        if (e != cudaSuccess) {
            cudaFree(tmp);
            return cuda_fail(e, "cudaMemcpyAsync(replies)");
        }
        d_replies = static_cast<const ptype_reply*>(tmp);
    }
    cudaError_t e = ptype::launch_reply_checksum(d_replies, n, index_base, d->d_cksum_partials,
                                                 d->d_cksum_out, d->stream);
    ctx->launches.fetch_add(2);
    unsigned long long out = 0;
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(&out, d->d_cksum_out, sizeof out, cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(d->stream);
    if (tmp) cudaFree(tmp);
    if (e != cudaSuccess) return cuda_fail(e, "reply_checksum");
    *checksum = out;
    return PTYPE_OK;
}

int ptype_b200_resolve_names(ptype_ctx* ctx, int dev, const char* names, uint32_t stride, size_t n,
                             ptype_method_ref* out, ptype_req* patch) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || (n && !names) || stride == 0 || (stride & 15u) || (!out && !patch)) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    d->after_dispatch = false;
    CU(ptype::launch_resolve_names(reinterpret_cast<const unsigned char*>(names), stride, n, out, patch,
                                   d->stream));
    ctx->launches.fetch_add(1);
    CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}

int ptype_b200_gob_encode_replies(ptype_ctx* ctx, int dev, const ptype_reply* replies, size_t n,
                                  uint64_t* offsets, uint8_t* out, size_t out_cap, uint64_t* total_bytes) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !offsets || !total_bytes || (n && !replies) || (out_cap && !out)) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    const size_t need = ptype::gob_scratch_bytes(n) + 64;
    if (need > d->codec_cap) {
        CU(cudaStreamSynchronize(d->stream));
        if (d->d_codec) CU(cudaFree(d->d_codec));
        d->d_codec = nullptr;
        d->codec_cap = 0;
        CU(cudaMalloc(&d->d_codec, need));
        d->codec_cap = need;
    }
    if (!d->d_cksum_out) CU(cudaMalloc(&d->d_cksum_out, sizeof(unsigned long long)));
    d->after_dispatch = false;
    CU(ptype::launch_gob_encode_replies(replies, n, reinterpret_cast<unsigned long long*>(offsets), out,
                                        out_cap, d->d_codec, d->d_cksum_out, d->stream));
    ctx->launches.fetch_add(3);
    unsigned long long total = 0;
    if (n) CU(cudaMemcpyAsync(&total, d->d_cksum_out, sizeof total, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    *total_bytes = total;
    return PTYPE_OK;
}

int ptype_b200_gob_decode_args(ptype_ctx* ctx, int dev, const uint8_t* bytes, const uint64_t* offsets, size_t n,
                               int64_t* args, int args_per_call, uint8_t* status) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || (n && (!bytes || !offsets || !args)) || (args_per_call != 2 && args_per_call != 3))
        return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    d->after_dispatch = false;
    CU(ptype::launch_gob_decode_args(bytes, reinterpret_cast<const unsigned long long*>(offsets), n, args,
                                     args_per_call, status, d->stream));
    ctx->launches.fetch_add(1);
    CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}

int ptype_b200_wire_decode(ptype_ctx* ctx, int dev, const uint8_t* bytes, const uint64_t* stream_off, size_t n_streams,
                           const ptype_wire_stream* streams, uint64_t* call_off, ptype_req* reqs,
                           ptype_wire_call* calls, size_t cap, uint64_t* total_calls) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !call_off || !total_calls || (n_streams && (!bytes || !stream_off || !streams))) return PTYPE_E_INVAL;
    if (misaligned(reqs, 16) || misaligned(calls, 8) || misaligned(streams, 8) || misaligned(bytes, 16))
        return PTYPE_E_INVAL;  // the decoder copies the streams into shared memory 16 bytes at a time
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    d->after_dispatch = false;
    unsigned long long* off = reinterpret_cast<unsigned long long*>(call_off);
    CU(ptype::launch_wire_count(bytes, reinterpret_cast<const unsigned long long*>(stream_off), n_streams, off,
                                d->stream));
    ctx->launches.fetch_add(2);
    unsigned long long total = 0;
    CU(cudaMemcpyAsync(&total, off + n_streams, sizeof total, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    *total_calls = total;
    if (total > cap || (total && (!reqs || !calls))) return PTYPE_E_LIMIT;
    CU(ptype::launch_wire_decode(bytes, reinterpret_cast<const unsigned long long*>(stream_off), n_streams, streams,
                                 off, reqs, calls, d->stream));
    ctx->launches.fetch_add(1);
    CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}

int ptype_b200_wire_encode(ptype_ctx* ctx, int dev, const uint8_t* request_bytes, const uint64_t* call_off,
                           size_t n_streams, const ptype_wire_call* calls, const ptype_reply* replies,
                           uint32_t response_id, uint32_t empty_id, uint8_t* state, uint64_t* out_off, uint8_t* out,
                           size_t out_cap, uint64_t* total_bytes) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !out_off || !total_bytes || (n_streams && (!call_off || !state))) return PTYPE_E_INVAL;
    if (response_id == 0 || empty_id == 0 || response_id > 0x3FFFFFFFu || empty_id > 0x3FFFFFFFu) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->device));
    // the two type definitions for these ids, kept on the device (rebuilt when the ids change)
    if (d->wire_ids[0] != response_id || d->wire_ids[1] != empty_id || !d->d_wire_defs) {
        std::string a, b;
        ptype::wire_server_typedefs(response_id, empty_id, &a, &b);
        const std::string both = a + b;
        CU(cudaStreamSynchronize(d->stream));
        if (!d->d_wire_defs) CU(cudaMalloc(&d->d_wire_defs, 512));
        if (both.size() > 512) return PTYPE_E_LIMIT;
        CU(cudaMemcpy(d->d_wire_defs, both.data(), both.size(), cudaMemcpyHostToDevice));
        d->wire_ids[0] = response_id;
        d->wire_ids[1] = empty_id;
        d->wire_def_len[0] = static_cast<uint32_t>(a.size());
        d->wire_def_len[1] = static_cast<uint32_t>(b.size());
    }
    d->after_dispatch = false;
    unsigned long long* off = reinterpret_cast<unsigned long long*>(out_off);
    const unsigned long long* coff = reinterpret_cast<const unsigned long long*>(call_off);
    CU(ptype::launch_wire_encode(true, request_bytes, coff, n_streams, calls, replies, d->d_wire_defs,
                                 d->wire_def_len[0], d->wire_def_len[1], response_id, empty_id, state, off, out,
                                 d->stream));
    ctx->launches.fetch_add(2);
    unsigned long long total = 0;
    CU(cudaMemcpyAsync(&total, off + n_streams, sizeof total, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    *total_bytes = total;
    if (total > out_cap || (total && !out)) return PTYPE_E_LIMIT;
    CU(ptype::launch_wire_encode(false, request_bytes, coff, n_streams, calls, replies, d->d_wire_defs,
                                 d->wire_def_len[0], d->wire_def_len[1], response_id, empty_id, state, off, out,
                                 d->stream));
    ctx->launches.fetch_add(1);
    CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}

namespace {
// Slot of a service in the host copy of the active image's table, or PTYPE_SLOT_EMPTY.
uint32_t host_slot_of(const ptype_ctx* ctx, const char* service) {
    const ptype_image_header* hh = reinterpret_cast<const ptype_image_header*>(ctx->host_image.data());
    const uint64_t key = ptype_b200_service_key(service, hh->key_salt);
    const ptype_slot* hs = reinterpret_cast<const ptype_slot*>(ctx->host_image.data() + hh->off_slots);
    for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
        const uint32_t c = ptype::slot_of(static_cast<uint32_t>(key), static_cast<uint32_t>(key >> 32),
                                          hh->table_seed, w, hh->n_slots);
        if (hs[c].node_base != PTYPE_SLOT_EMPTY && hs[c].key == key) return c;
    }
    return PTYPE_SLOT_EMPTY;
}
}  // namespace

namespace {
// Shared body of assign_seq (numbers into the records) and assign_seq_array (into d_seq).
int assign_seq_impl(ptype_ctx* ctx, int dev, ptype_req* d_reqs, size_t n, unsigned long long* d_seq, uint32_t flags) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || (n && !d_reqs) || misaligned(d_reqs, 16)) return PTYPE_E_INVAL;
    if (n > 0xFFFFFFFFull) return PTYPE_E_LIMIT;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    CU(cudaSetDevice(d->device));
    const ptype::DeviceImage img = d->view[d->active];
    if (!ptype::rank_supported(img.n_slots)) {
        g_last_error = "ptype_b200_assign_seq: the per-CTA histogram holds at most 51200 table slots";
        return PTYPE_E_LIMIT;
    }
    int rc = ensure_counters(d, d->active, img.n_slots);
    if (rc != PTYPE_OK) return rc;
    d->seq_in_use = true;
    if (n == 0) return PTYPE_OK;
    ptype::RankPlan plan;
    CU(ptype::rank_plan(img.n_slots, n, d->dispatch_sms, &plan));
    const size_t cells = static_cast<size_t>(plan.n_chunks) * img.n_slots;
    if (cells > d->rank_rows_cap) {
        CU(cudaStreamSynchronize(d->stream));
        if (d->rank_rows) CU(cudaFree(d->rank_rows));
        d->rank_rows = nullptr;
        d->rank_rows_cap = 0;
        CU(cudaMalloc(&d->rank_rows, cells * sizeof(uint32_t)));
        d->rank_rows_cap = cells;
    }
    if (n > d->rank_ids_cap) {
        CU(cudaStreamSynchronize(d->stream));
        if (d->rank_ids) CU(cudaFree(d->rank_ids));
        d->rank_ids = nullptr;
        d->rank_ids_cap = 0;
        CU(cudaMalloc(&d->rank_ids, n * sizeof(uint32_t)));
        d->rank_ids_cap = n;
    }
    if (img.n_slots > d->seq_start_cap) {
        CU(cudaStreamSynchronize(d->stream));
        if (d->seq_start) CU(cudaFree(d->seq_start));
        d->seq_start = nullptr;
        d->seq_start_cap = 0;
        CU(cudaMalloc(&d->seq_start, static_cast<size_t>(img.n_slots) * sizeof(unsigned long long)));
        d->seq_start_cap = img.n_slots;
    }
    d->after_dispatch = false;
    CU(ptype::launch_assign_seq(img, plan, d_reqs, n, d->rank_rows, d->rank_ids, d->seq_counters[d->active],
                                d->seq_start, d_seq, d->stream));
    ctx->launches.fetch_add(3);
    if (!(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}
}  // namespace

int ptype_b200_assign_seq(ptype_ctx* ctx, int dev, ptype_req* d_reqs, size_t n, uint32_t flags) {
    return assign_seq_impl(ctx, dev, d_reqs, n, nullptr, flags);
}

int ptype_b200_assign_seq_array(ptype_ctx* ctx, int dev, const ptype_req* d_reqs, size_t n, uint64_t* d_seq,
                                uint32_t flags) {
    if ((n && !d_seq) || misaligned(d_seq, 8)) return PTYPE_E_INVAL;
    if (n == 0) return assign_seq_impl(ctx, dev, nullptr, 0, nullptr, flags);
    // the records are only read when the numbers go to an array
    return assign_seq_impl(ctx, dev, const_cast<ptype_req*>(d_reqs), n, reinterpret_cast<unsigned long long*>(d_seq),
                           flags);
}

int ptype_b200_dispatch_seq(ptype_ctx* ctx, int dev, const ptype_req* d_reqs, const uint64_t* d_seq, size_t n,
                            ptype_reply* d_replies, uint32_t flags) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || (n && (!d_reqs || !d_seq || !d_replies))) return PTYPE_E_INVAL;
    if (misaligned(d_reqs, 16) || misaligned(d_replies, 16) || misaligned(d_seq, 8)) return PTYPE_E_INVAL;
    if (n > 0xFFFFFFFFull) return PTYPE_E_LIMIT;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    CU(cudaSetDevice(d->device));
    const int buf = d->active;
    const ptype::DeviceImage img = d->view[buf];
    // the loop that reads the side array loads requests with LDG: that plan's block shape
    const ptype::LaunchPlan& plan = d->plan[buf][plan_kind((flags & PTYPE_F_GLOBAL_TABLE) | PTYPE_F_LDG_LOADS)][0];
    if (n) {
        const ptype_batch one = {d_reqs, d_replies, n};
        // d_seq was written by whatever ran before on this stream (assign_seq_array): no overlap with it
        d->after_dispatch = false;
        CU(ptype::launch_dispatch_multi(img, plan, d->dispatch_sms, &one, 1, d->stream, take_pdl(d, flags),
                                        reinterpret_cast<const unsigned long long*>(d_seq)));
        ctx->launches.fetch_add(1);
    }
    if (!(flags & PTYPE_F_ASYNC)) CU(cudaStreamSynchronize(d->stream));
    return PTYPE_OK;
}

int ptype_b200_seq_get(ptype_ctx* ctx, int dev, const char* service, uint64_t* value) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !service || !value) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> ilk(ctx->image_mu);
    if (ctx->host_image.empty()) return PTYPE_E_NO_IMAGE;
    const uint32_t at = host_slot_of(ctx, service);
    if (at == PTYPE_SLOT_EMPTY) return PTYPE_E_NOT_FOUND;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    *value = 0;
    if (!d->seq_in_use) return PTYPE_OK;
    CU(cudaSetDevice(d->device));
    CU(cudaStreamSynchronize(d->stream));
    unsigned long long v = 0;
    CU(cudaMemcpy(&v, d->seq_counters[d->active] + at, sizeof v, cudaMemcpyDeviceToHost));
    *value = v;
    return PTYPE_OK;
}

int ptype_b200_seq_set(ptype_ctx* ctx, int dev, const char* service, uint64_t value) {
    DeviceState* d = dev_of(ctx, dev);
    if (!d || !service) return PTYPE_E_INVAL;
    std::lock_guard<std::mutex> ilk(ctx->image_mu);
    if (ctx->host_image.empty()) return PTYPE_E_NO_IMAGE;
    const uint32_t at = host_slot_of(ctx, service);
    if (at == PTYPE_SLOT_EMPTY) return PTYPE_E_NOT_FOUND;
    std::lock_guard<std::mutex> lk(d->mu);
    if (d->active < 0) return PTYPE_E_NO_IMAGE;
    CU(cudaSetDevice(d->device));
    int rc = ensure_counters(d, d->active, d->view[d->active].n_slots);
    if (rc != PTYPE_OK) return rc;
    d->seq_in_use = true;
    CU(cudaStreamSynchronize(d->stream));
    const unsigned long long v = value;
    CU(cudaMemcpy(d->seq_counters[d->active] + at, &v, sizeof v, cudaMemcpyHostToDevice));
    return PTYPE_OK;
}

int ptype_b200_launch_count(ptype_ctx* ctx, uint64_t* launches) {
    if (!ctx || !launches) return PTYPE_E_INVAL;
    *launches = ctx->launches.load();
    return PTYPE_OK;
}

}  // extern "C"
