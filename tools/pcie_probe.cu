// pcie_probe — how fast can kernels stream pinned HOST memory over PCIe, and with which access shape?
// Background for the zero-copy host-buffer path of ptype_b200_call_batch[_values]: the kernel reads
// its arguments from, and writes its replies to, mapped pinned memory.  This probe times, with CUDA
// events on the launching stream:
//   copy engines (cudaMemcpyAsync) H2D, D2H and both at once,
//   kernel reads with 128-bit loads (the current shape), at several grid sizes and unroll depths,
//   kernel reads with bulk asynchronous copies (TMA, 1-D) of 2/4/8/16 KiB per request,
//   kernel writes with 128-bit and 64-bit stores,
//   the dispatch mix: 16 B in + 8 B out per unit, loads vs bulk copies.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pcie_probe pcie_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <chrono>
#include <string>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            std::exit(1);                                                                      \
        }                                                                                      \
    } while (0)

__device__ __forceinline__ uint4 ld_nc(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

template <int kUnroll>
__global__ void __launch_bounds__(256) read_ld128(const uint4* __restrict__ src, size_t n16,
                                                  unsigned long long* __restrict__ sink) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    unsigned long long acc = 0;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n16;
         i += stride * kUnroll) {
        uint4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t j = i + u * stride;
            v[u] = j < n16 ? ld_nc(src + j) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x123456789abcdefull) sink[0] = acc;
}

// One warp = one ring of kDepth chunks of `chunk` bytes, filled by bulk copies from host memory.
template <int kDepth>
__global__ void __launch_bounds__(256) read_bulk(const unsigned char* __restrict__ src, size_t bytes,
                                                 uint32_t chunk, unsigned long long* __restrict__ sink) {
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t bars[8][kDepth];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t warps = blockDim.x >> 5;
    unsigned char* my = ring + static_cast<size_t>(warp) * kDepth * chunk;
    if (lane == 0)
        for (int d = 0; d < kDepth; ++d) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][d])));
        }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const size_t n_chunks = bytes / chunk;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * warps;
    const size_t w = static_cast<size_t>(blockIdx.x) * warps + warp;
    auto issue = [&](size_t c, int d) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[warp][d])),
                     "r"(chunk)
                     : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                smem_u32(my + static_cast<size_t>(d) * chunk)),
            "l"(src + c * chunk), "r"(chunk), "r"(smem_u32(&bars[warp][d]))
            : "memory");
    };
    if (lane == 0)
        for (int d = 0; d < kDepth; ++d) {
            const size_t c = w + d * total_warps;
            if (c < n_chunks) issue(c, d);
        }
    unsigned long long acc = 0;
    uint32_t k = 0;
    for (size_t c = w; c < n_chunks; c += total_warps, ++k) {
        const int d = k % kDepth;
        const uint32_t parity = (k / kDepth) & 1;
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                : "=r"(done)
                : "r"(smem_u32(&bars[warp][d])), "r"(parity)
                : "memory");
        }
        acc += reinterpret_cast<const uint32_t*>(my + static_cast<size_t>(d) * chunk)[lane];
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        const size_t nxt = c + static_cast<size_t>(kDepth) * total_warps;
        if (lane == 0 && nxt < n_chunks) issue(nxt, d);
    }
    if (acc == 0x123456789abcdefull) sink[0] = acc;
}

template <int kBytes>
__global__ void __launch_bounds__(256) write_st(unsigned char* __restrict__ dst, size_t n_units) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_units; i += stride) {
        if (kBytes == 16) {
            const uint32_t x = static_cast<uint32_t>(i);
            asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + i * 16), "r"(x), "r"(x), "r"(x),
                         "r"(x)
                         : "memory");
        } else {
            reinterpret_cast<unsigned long long*>(dst)[i] = i;
        }
    }
}

// The dispatch mix with loads/stores: 16 B in, 8 B out per unit.
template <int kUnroll>
__global__ void __launch_bounds__(256) mix_ld(const uint4* __restrict__ src, unsigned long long* __restrict__ dst,
                                              size_t n) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride * kUnroll) {
        uint4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t j = i + u * stride;
            v[u] = j < n ? ld_nc(src + j) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t j = i + u * stride;
            if (j < n)
                dst[j] = (static_cast<unsigned long long>(v[u].x) | (static_cast<unsigned long long>(v[u].y) << 32)) *
                         (static_cast<unsigned long long>(v[u].z) | (static_cast<unsigned long long>(v[u].w) << 32));
        }
    }
}

// The dispatch mix with bulk copies in (chunk bytes = chunk/16 units) and plain 8-byte stores out.
template <int kDepth>
__global__ void __launch_bounds__(256) mix_bulk(const unsigned char* __restrict__ src,
                                                unsigned long long* __restrict__ dst, size_t n, uint32_t chunk) {
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t bars[8][kDepth];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t warps = blockDim.x >> 5;
    unsigned char* my = ring + static_cast<size_t>(warp) * kDepth * chunk;
    if (lane == 0)
        for (int d = 0; d < kDepth; ++d)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][d])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const size_t n_chunks = n * 16 / chunk;
    const uint32_t units = chunk / 16;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * warps;
    const size_t w = static_cast<size_t>(blockIdx.x) * warps + warp;
    auto issue = [&](size_t c, int d) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[warp][d])),
                     "r"(chunk)
                     : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                smem_u32(my + static_cast<size_t>(d) * chunk)),
            "l"(src + c * chunk), "r"(chunk), "r"(smem_u32(&bars[warp][d]))
            : "memory");
    };
    if (lane == 0)
        for (int d = 0; d < kDepth; ++d) {
            const size_t c = w + d * total_warps;
            if (c < n_chunks) issue(c, d);
        }
    uint32_t k = 0;
    for (size_t c = w; c < n_chunks; c += total_warps, ++k) {
        const int d = k % kDepth;
        const uint32_t parity = (k / kDepth) & 1;
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                : "=r"(done)
                : "r"(smem_u32(&bars[warp][d])), "r"(parity)
                : "memory");
        }
        const uint4* stage = reinterpret_cast<const uint4*>(my + static_cast<size_t>(d) * chunk);
        for (uint32_t u = lane; u < units; u += 32) {
            const uint4 v = stage[u];
            dst[c * units + u] = (static_cast<unsigned long long>(v.x) | (static_cast<unsigned long long>(v.y) << 32)) *
                                 (static_cast<unsigned long long>(v.z) | (static_cast<unsigned long long>(v.w) << 32));
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        const size_t nxt = c + static_cast<size_t>(kDepth) * total_warps;
        if (lane == 0 && nxt < n_chunks) issue(nxt, d);
    }
}

static cudaStream_t g_stream;

template <class F>
static double time_ms(F&& launch, int reps = 7) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    std::vector<float> t;
    for (int r = 0; r < reps + 2; ++r) {
        CK(cudaEventRecord(a, g_stream));
        launch();
        CK(cudaEventRecord(b, g_stream));
        CK(cudaEventSynchronize(b));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (r >= 2) t.push_back(ms);
    }
    std::sort(t.begin(), t.end());
    CK(cudaEventDestroy(a));
    CK(cudaEventDestroy(b));
    return t[t.size() / 2];
}

static void report(const char* what, double bytes, double ms) {
    std::printf("%-64s %8.3f ms  %7.2f GB/s\n", what, ms, bytes / ms * 1e-6);
    std::fflush(stdout);
}

// `pcie_probe concurrent <device> <seconds> [wc]`: the box ceiling when several GPUs stream host memory
// at once.  Start one process per GPU at the same moment; each keeps ONE shape running for `seconds`
// and prints its sustained rate: the product's mix (16 B in / 8 B out per unit, kernel loads and
// stores on mapped pinned memory, one CTA per SM), then the two copy engines at once, then the
// hybrid (copy engine in, kernel stores out).  `wc` allocates the input write-combined.
static int run_concurrent(int argc, char** argv) {
    const int device = argc > 2 ? std::atoi(argv[2]) : 0;
    const double seconds = argc > 3 ? std::atof(argv[3]) : 2.0;
    const bool wc = argc > 4 && std::string(argv[4]) == "wc";
    const size_t n = 1u << 20, in_bytes = n * 16, out_bytes = n * 8;
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
    cudaStream_t s2;
    CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    unsigned char *h_in, *h_out, *d_in, *d_out;
    CK(cudaHostAlloc(reinterpret_cast<void**>(&h_in), in_bytes,
                     cudaHostAllocMapped | cudaHostAllocPortable | (wc ? cudaHostAllocWriteCombined : 0)));
    CK(cudaHostAlloc(reinterpret_cast<void**>(&h_out), in_bytes, cudaHostAllocMapped | cudaHostAllocPortable));
    CK(cudaMalloc(reinterpret_cast<void**>(&d_in), in_bytes));
    CK(cudaMalloc(reinterpret_cast<void**>(&d_out), in_bytes));
    for (size_t i = 0; i < in_bytes; i += 8) *reinterpret_cast<unsigned long long*>(h_in + i) = i * 0x9E3779B97F4A7C15ull;
    cudaEvent_t fork, join;
    CK(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&join, cudaEventDisableTiming));
    auto sustained = [&](const char* what, auto&& one) {
        const auto t0 = std::chrono::steady_clock::now();
        size_t reps = 0;
        double el = 0;
        do {
            for (int k = 0; k < 8; ++k) one();
            CK(cudaStreamSynchronize(g_stream));
            reps += 8;
            el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        } while (el < seconds);
        std::printf("gpu %d%s %-58s %8.1f us/batch %7.2f GB/s %6.2f G units/s\n", device, wc ? " wc" : "", what,
                    el / reps * 1e6, (in_bytes + out_bytes) * reps / el * 1e-9, n * reps / el * 1e-9);
        std::fflush(stdout);
    };
    sustained("mix: kernel ld.128 in + st.64 out on host memory, 148 CTAs", [&] {
        mix_ld<2><<<148, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(h_in), reinterpret_cast<unsigned long long*>(h_out), n);
    });
    sustained("copy engines: H2D in + D2H out at once", [&] {
        CK(cudaEventRecord(fork, g_stream));
        CK(cudaStreamWaitEvent(s2, fork, 0));
        CK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, g_stream));
        CK(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, s2));
        CK(cudaEventRecord(join, s2));
        CK(cudaStreamWaitEvent(g_stream, join, 0));
    });
    sustained("hybrid: copy engine H2D, kernel HBM in / st.64 host out", [&] {
        CK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, g_stream));
        mix_ld<2><<<148 * 4, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(d_in), reinterpret_cast<unsigned long long*>(h_out), n);
    });
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::string(argv[1]) == "concurrent") return run_concurrent(argc, argv);
    const size_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 0) : (1u << 20);  // units (calls)
    const size_t in_bytes = n * 16, out_bytes = n * 8;
    CK(cudaSetDevice(0));
    CK(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
    unsigned char *h_in, *h_out, *d_in, *d_out;
    unsigned long long* d_sink;
    CK(cudaHostAlloc(reinterpret_cast<void**>(&h_in), in_bytes, cudaHostAllocMapped | cudaHostAllocPortable));
    CK(cudaHostAlloc(reinterpret_cast<void**>(&h_out), in_bytes, cudaHostAllocMapped | cudaHostAllocPortable));
    CK(cudaMalloc(reinterpret_cast<void**>(&d_in), in_bytes));
    CK(cudaMalloc(reinterpret_cast<void**>(&d_out), in_bytes));
    CK(cudaMalloc(reinterpret_cast<void**>(&d_sink), 64));
    for (size_t i = 0; i < in_bytes; ++i) h_in[i] = static_cast<unsigned char>(i * 2654435761u >> 13);
    std::printf("units %zu: %.1f MB in (16 B each), %.1f MB out (8 B each)\n", n, in_bytes * 1e-6, out_bytes * 1e-6);

    cudaStream_t s2;
    CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    cudaEvent_t fork, join;
    CK(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&join, cudaEventDisableTiming));
    report("copy engine H2D (in bytes)", in_bytes,
           time_ms([&] { CK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, g_stream)); }));
    report("copy engine D2H (out bytes)", out_bytes,
           time_ms([&] { CK(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, g_stream)); }));
    report("copy engines H2D in + D2H out at once (in+out bytes)", in_bytes + out_bytes, time_ms([&] {
               CK(cudaEventRecord(fork, g_stream));
               CK(cudaStreamWaitEvent(s2, fork, 0));
               CK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, g_stream));
               CK(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, s2));
               CK(cudaEventRecord(join, s2));
               CK(cudaStreamWaitEvent(g_stream, join, 0));
           }));

    char label[128];
    for (int ctas : {1, 2, 4, 8}) {
        const int grid = 148 * ctas;
        std::snprintf(label, sizeof label, "kernel read ld.128, grid 148x%d, unroll 1", ctas);
        report(label, in_bytes, time_ms([&] {
                   read_ld128<1><<<grid, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(h_in), n, d_sink);
               }));
        std::snprintf(label, sizeof label, "kernel read ld.128, grid 148x%d, unroll 4", ctas);
        report(label, in_bytes, time_ms([&] {
                   read_ld128<4><<<grid, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(h_in), n, d_sink);
               }));
        std::snprintf(label, sizeof label, "kernel read ld.128, grid 148x%d, unroll 8", ctas);
        report(label, in_bytes, time_ms([&] {
                   read_ld128<8><<<grid, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(h_in), n, d_sink);
               }));
    }
    CK(cudaFuncSetAttribute(read_bulk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(read_bulk<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(mix_bulk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(mix_bulk<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (uint32_t chunk : {2048u, 4096u, 8192u, 16384u}) {
        for (int threads : {64, 256}) {
            const int warps = threads / 32;
            for (int depth : {2, 4}) {
                const size_t smem = static_cast<size_t>(warps) * depth * chunk;
                if (smem > 200 * 1024) continue;
                std::snprintf(label, sizeof label, "kernel read bulk %5u B x depth %d, grid 148 x %d threads", chunk, depth,
                              threads);
                report(label, in_bytes, time_ms([&] {
                           if (depth == 2)
                               read_bulk<2><<<148, threads, smem, g_stream>>>(h_in, in_bytes, chunk, d_sink);
                           else
                               read_bulk<4><<<148, threads, smem, g_stream>>>(h_in, in_bytes, chunk, d_sink);
                       }));
            }
        }
    }
    for (int ctas : {1, 4}) {
        std::snprintf(label, sizeof label, "kernel write st.128 to host, grid 148x%d (in-size bytes)", ctas);
        report(label, in_bytes, time_ms([&] { write_st<16><<<148 * ctas, 256, 0, g_stream>>>(h_out, n); }));
        std::snprintf(label, sizeof label, "kernel write st.64 to host, grid 148x%d (out bytes)", ctas);
        report(label, out_bytes, time_ms([&] { write_st<8><<<148 * ctas, 256, 0, g_stream>>>(h_out, n); }));
    }
    for (int ctas : {1, 2, 4, 8}) {
        std::snprintf(label, sizeof label, "mix 16 B in / 8 B out, ld.128 + st.64, grid 148x%d, unroll 2 (in+out)", ctas);
        report(label, in_bytes + out_bytes, time_ms([&] {
                   mix_ld<2><<<148 * ctas, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(h_in),
                                                              reinterpret_cast<unsigned long long*>(h_out), n);
               }));
        std::snprintf(label, sizeof label, "mix 16 B in / 8 B out, ld.128 + st.64, grid 148x%d, unroll 8 (in+out)", ctas);
        report(label, in_bytes + out_bytes, time_ms([&] {
                   mix_ld<8><<<148 * ctas, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(h_in),
                                                              reinterpret_cast<unsigned long long*>(h_out), n);
               }));
    }
    for (uint32_t chunk : {4096u, 16384u}) {
        for (int depth : {2, 4}) {
            const int threads = 256;
            const size_t smem = static_cast<size_t>(threads / 32) * depth * chunk;
            if (smem > 200 * 1024) continue;
            std::snprintf(label, sizeof label, "mix 16 B in / 8 B out, bulk %5u B x depth %d in + st.64 out (in+out)", chunk,
                          depth);
            report(label, in_bytes + out_bytes, time_ms([&] {
                       if (depth == 2)
                           mix_bulk<2><<<148, threads, smem, g_stream>>>(h_in, reinterpret_cast<unsigned long long*>(h_out), n,
                                                                        chunk);
                       else
                           mix_bulk<4><<<148, threads, smem, g_stream>>>(h_in, reinterpret_cast<unsigned long long*>(h_out), n,
                                                                        chunk);
                   }));
        }
    }
    // in over the copy engine into HBM, out by kernel stores to host: the hybrid
    report("hybrid: copy engine H2D, then mix kernel HBM in / host out (in+out)", in_bytes + out_bytes, time_ms([&] {
               CK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, g_stream));
               mix_ld<2><<<148 * 4, 256, 0, g_stream>>>(reinterpret_cast<const uint4*>(d_in),
                                                        reinterpret_cast<unsigned long long*>(h_out), n);
           }));
    return 0;
}
