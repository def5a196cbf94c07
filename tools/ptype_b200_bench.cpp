// ptype_b200_bench.cpp — the C ABI driven from plain C++ (no Python, no torch): builds the
// BASELINE config-2 registry and call stream, keeps `buffers` batches resident in HBM, submits
// them with ptype_b200_dispatch_batches and times the stream with CUDA events.
//
//   ./ptype_b200_bench [calls_per_batch=1048576] [steps=2000] [buffers=8] [mesh=0]
//
// Prints calls/s, microseconds per batch and the order-independent checksum of batch 0's
// replies (tests/test_tools_gpu.py compares it with the oracle's for the same stream).
//
//   ./ptype_b200_bench mirror [calls_per_batch=1048576] [steps=200] [device=0] [mesh=0]
//
// End to end through the C++ mirror of the reference's client (ptype_b200/host/cluster.hpp):
// a MemRegistry holding the config-2 registrations, a Dispatcher, one Client created through
// the registry watch, and `steps` x Client::CallBatch on the client's page-locked Batch - the
// sequence reservation, the C-ABI call, the PCIe transfers and the failed-list read are all
// inside the timing.  Then the same batch through the array form of CallBatch (arguments copied
// in, 16-byte reply records unpacked).  Values are checked against plain C++ arithmetic.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/ptype_b200.h"
#include "../ptype_b200/host/cluster.hpp"

#define CHECK_ABI(call)                                                                       \
    do {                                                                                      \
        int rc_ = (call);                                                                     \
        if (rc_ != PTYPE_OK) {                                                                \
            std::fprintf(stderr, "%s: %s (%s)\n", #call, ptype_b200_strerror(rc_), ptype_b200_last_error()); \
            return 1;                                                                         \
        }                                                                                     \
    } while (0)

// splitmix64 exactly as ptype_b200/workloads.py draws it: element i of stream s started at seed.
static uint64_t splitmix(uint64_t seed, uint64_t stream, uint64_t i) {
    uint64_t z = seed + stream * 0x632BE59BD9B4E019ull + (i + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// workloads.calculator_calls: 50/50 Add/Multiply, half the args full range, half in [-2^20, 2^20].
static void calculator_calls(std::vector<ptype_req>& out, size_t n, uint64_t seed, uint64_t svc_key) {
    uint32_t mul_m, mul_r, add_m, add_r;
    ptype_b200_method_key("Calculator.Multiply", &mul_m, &mul_r);
    ptype_b200_method_key("Calculator.Add", &add_m, &add_r);
    out.resize(n);
    const uint64_t span = (1ull << 21) + 1;
    for (size_t i = 0; i < n; ++i) {
        const uint64_t bits = splitmix(seed, 0, i);
        uint64_t a = splitmix(seed, 1, i), b = splitmix(seed, 2, i);
        if ((bits >> 1) & 1) {
            a = static_cast<uint64_t>(static_cast<int64_t>(a % span) - (1ll << 20));
            b = static_cast<uint64_t>(static_cast<int64_t>(b % span) - (1ll << 20));
        }
        ptype_req& q = out[i];
        q.svc_key = svc_key;
        q.seq = i + 1;
        q.method = (bits & 1) ? mul_m : add_m;
        q.recv = mul_r;
        q.a0 = static_cast<int64_t>(a);
        q.a1 = static_cast<int64_t>(b);
        q.a2 = 0;
    }
}

static int run_mirror(int argc, char** argv) {
    using namespace cluster;
    const size_t n = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : (1u << 20);
    const int steps = argc > 3 ? std::atoi(argv[3]) : 200;
    const int device = argc > 4 ? std::atoi(argv[4]) : 0;
    const int mesh = argc > 5 ? std::atoi(argv[5]) : 0;

    MemRegistry reg;   // what 16 calculator servers wrote through Registry.Register (registry.go:51-86)
    for (int i = 1; i <= 16; ++i) {
        char name[64], addr[64];
        std::snprintf(name, sizeof name, "calculator_node_%02d", i);
        std::snprintf(addr, sizeof addr, "10.0.0.%d", i);
        if (reg.Register("calculator", name, addr, 31234)) return 1;
    }
    Dispatcher gpu("10.0.0.5", mesh ? 0 : 3, device);
    if (gpu.Err()) {
        std::fprintf(stderr, "%s\n", gpu.Err().msg.c_str());
        return 1;
    }
    gpu.Expose("calculator", "Calculator");
    ConnConfig cfg = DefaultConnConfig();
    cfg.MaxConnections = mesh ? 0 : 3;
    std::unique_ptr<Client> client;
    if (Error e = Client::New(&gpu, "calculator", &reg, &cfg, &client)) {
        std::fprintf(stderr, "%s\n", e.msg.c_str());
        return 1;
    }
    // the config-2 stream, written straight into the client's page-locked batch
    std::vector<ptype_req> host;
    calculator_calls(host, n, 0x5EED0100ull, 0);
    uint32_t mul_m, mul_r;
    ptype_b200_method_key("Calculator.Multiply", &mul_m, &mul_r);
    auto batch = client->NewBatch(n, 2);
    if (!batch->Ok()) return 1;
    for (size_t j = 0; j < n; ++j) {
        batch->Args()[2 * j] = host[j].a0;
        batch->Args()[2 * j + 1] = host[j].a1;
        batch->Sel()[j] = host[j].method == mul_m;
    }
    const std::vector<std::string> methods = {"Calculator.Add", "Calculator.Multiply"};
    auto check = [&](const int64_t* values) {
        for (size_t j = 0; j < n; ++j) {
            const uint64_t a = static_cast<uint64_t>(host[j].a0), b = static_cast<uint64_t>(host[j].a1);
            if (values[j] != static_cast<int64_t>(batch->Sel()[j] ? a * b : a + b)) return false;
        }
        return true;
    };
    for (int i = 0; i < 3; ++i)
        if (Error e = client->CallBatch(methods, *batch, n)) {
            std::fprintf(stderr, "%s\n", e.msg.c_str());
            return 1;
        }
    bool ok = batch->Failed().empty() && check(batch->Values());
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; ++i)
        if (client->CallBatch(methods, *batch, n)) return 1;
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    ok = ok && batch->Failed().empty() && check(batch->Values()) && client->Seq() == static_cast<uint64_t>(steps + 3) * n;

    // the array form: one method per call, [n][3] arguments copied in, BatchReply records out
    std::vector<std::array<int64_t, 3>> plain(n);
    for (size_t j = 0; j < n; ++j) plain[j] = {host[j].a0, host[j].a1, 0};
    std::vector<BatchReply> reps(n);
    const int plain_steps = steps < 20 ? steps : 20;
    const auto args3 = reinterpret_cast<const int64_t(*)[3]>(plain.data());
    if (client->CallBatch("Calculator.Multiply", args3, n, reps.data())) return 1;
    const auto p0 = std::chrono::steady_clock::now();
    for (int i = 0; i < plain_steps; ++i)
        if (client->CallBatch("Calculator.Multiply", args3, n, reps.data())) return 1;
    const double plain_secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - p0).count();
    for (size_t j = 0; j < n && ok; ++j)
        ok = reps[j].Status == 0 && reps[j].Value == static_cast<int64_t>(static_cast<uint64_t>(host[j].a0) *
                                                                         static_cast<uint64_t>(host[j].a1));
    std::printf("calls %.0f seconds %.6f parity %s h2d_bytes_per_step %zu d2h_bytes_per_step %zu "
                "array_calls %.0f array_seconds %.6f\n",
                static_cast<double>(n) * steps, secs, ok ? "ok" : "WRONG", n * 17, n * 8,
                static_cast<double>(n) * plain_steps, plain_secs);
    return ok ? 0 : 2;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::string(argv[1]) == "mirror") return run_mirror(argc, argv);
    const size_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : (1u << 20);
    const int steps = argc > 2 ? std::atoi(argv[2]) : 2000;
    const int buffers = argc > 3 ? std::atoi(argv[3]) : 8;
    const int mesh = argc > 4 ? std::atoi(argv[4]) : 0;

    // registry: 1 service `calculator` x 16 nodes, Client on 10.0.0.5, MaxConnections 3 (or mesh)
    ptype_builder* b = nullptr;
    CHECK_ABI(ptype_b200_builder_new("10.0.0.5", mesh ? 0 : 3, &b));
    for (int i = 1; i <= 16; ++i) {
        char name[64], addr[64];
        std::snprintf(name, sizeof name, "calculator_node_%02d", i);
        std::snprintf(addr, sizeof addr, "10.0.0.%d", i);
        CHECK_ABI(ptype_b200_builder_register(b, "calculator", name, addr, 31234));
    }
    CHECK_ABI(ptype_b200_builder_expose(b, "calculator", "Calculator"));
    void* image = nullptr;
    size_t image_bytes = 0;
    CHECK_ABI(ptype_b200_builder_finish(b, 1, &image, &image_bytes));
    ptype_b200_builder_free(b);
    const uint64_t salt = static_cast<const ptype_image_header*>(image)->key_salt;

    int device = 0;
    ptype_ctx* ctx = nullptr;
    CHECK_ABI(ptype_b200_open(&device, 1, &ctx));
    CHECK_ABI(ptype_b200_publish_image(ctx, image, image_bytes));
    ptype_b200_image_free(image);

    std::vector<ptype_batch> ring(buffers);
    std::vector<ptype_req> host;
    for (int k = 0; k < buffers; ++k) {
        calculator_calls(host, n, 0x5EED0002ull + k, ptype_b200_service_key("calculator", salt));
        void *d_req = nullptr, *d_rep = nullptr;
        CHECK_ABI(ptype_b200_device_alloc(ctx, 0, n * sizeof(ptype_req), &d_req));
        CHECK_ABI(ptype_b200_device_alloc(ctx, 0, n * sizeof(ptype_reply), &d_rep));
        CHECK_ABI(ptype_b200_memcpy_h2d(ctx, 0, d_req, host.data(), n * sizeof(ptype_req)));
        ring[k] = ptype_batch{static_cast<const ptype_req*>(d_req), static_cast<ptype_reply*>(d_rep), n};
    }
    std::vector<ptype_batch> timed(steps);
    for (int i = 0; i < steps; ++i) timed[i] = ring[i % buffers];

    void* stream_ptr = nullptr;
    CHECK_ABI(ptype_b200_stream(ctx, 0, &stream_ptr));
    cudaStream_t stream = static_cast<cudaStream_t>(stream_ptr);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    CHECK_ABI(ptype_b200_dispatch_batches(ctx, 0, ring.data(), ring.size(), PTYPE_F_DEVICE_PTRS));  // warm-up
    // PTYPE_BENCH_REPS=r: time the same submission r times, each on its own (synchronised before and
    // after), and report the fastest; the per-repetition times go to stderr
    const char* reps_env = std::getenv("PTYPE_BENCH_REPS");
    const int reps = reps_env ? std::atoi(reps_env) : 1;
    float ms = 0;
    // PTYPE_BENCH_PREROLL=p: batches kept in flight right before the timed submission (default: as
    // many as are timed), so that the timed one meets a GPU that has been busy for a while
    const char* pre_env = std::getenv("PTYPE_BENCH_PREROLL");
    std::vector<ptype_batch> preroll(pre_env ? std::atoi(pre_env) : steps);
    for (size_t i = 0; i < preroll.size(); ++i) preroll[i] = ring[i % buffers];
    for (int rep = 0; rep < (reps > 0 ? reps : 1); ++rep) {
        CHECK_ABI(ptype_b200_dispatch_batches(ctx, 0, preroll.data(), preroll.size(), PTYPE_F_DEVICE_PTRS | PTYPE_F_ASYNC));
        cudaEventRecord(e0, stream);
        CHECK_ABI(ptype_b200_dispatch_batches(ctx, 0, timed.data(), timed.size(), PTYPE_F_DEVICE_PTRS | PTYPE_F_ASYNC));
        cudaEventRecord(e1, stream);
        CHECK_ABI(ptype_b200_sync(ctx, 0));
        float one = 0;
        cudaEventElapsedTime(&one, e0, e1);
        if (reps > 1) std::fprintf(stderr, "rep %d: %.1f us\n", rep, one * 1e3);
        if (rep == 0 || one < ms) ms = one;
    }

    uint64_t checksum = 0;
    CHECK_ABI(ptype_b200_reply_checksum(ctx, 0, ring[0].replies, n, 0, &checksum, PTYPE_F_DEVICE_PTRS));
    const double calls = static_cast<double>(n) * steps;
    std::printf("calls_per_sec %.6e us_per_batch %.3f hbm_gbs %.1f checksum %llu\n", calls / (ms * 1e-3),
                ms * 1e3 / steps, calls * 64.0 / (ms * 1e-3) / 1e9, static_cast<unsigned long long>(checksum));
    for (auto& bt : ring) {
        ptype_b200_device_free(ctx, 0, const_cast<ptype_req*>(bt.reqs));
        ptype_b200_device_free(ctx, 0, bt.replies);
    }
    ptype_b200_close(ctx);
    return 0;
}
