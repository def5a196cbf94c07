// spread_probe.cu — how unevenly do the SMs of a B200 finish a statically partitioned HBM stream?
//
// The dispatch kernels give every warp the same number of 32-call tiles.  This probe runs a
// kernel of the same shape (persistent grid, 4 CTAs of 256 threads per SM, three coalesced
// 128-bit loads and one 128-bit store per lane and tile, round-robin tiles) over `batches`
// batches of 2^20 records and records, per CTA, the SM it ran on and the time its last tile was
// done.  Output: the finish-time spread, per SM and per group of SMs.
//
//   ./spread_probe [batches=20] [repeats=5]
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                      \
            return 1;                                                                          \
        }                                                                                      \
    } while (0)

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned smid() {
    unsigned s;
    asm volatile("mov.u32 %0, %smid;" : "=r"(s));
    return s;
}

struct Mark {
    unsigned long long start, end;
    unsigned sm, tiles;
};

__global__ void __launch_bounds__(256, 4)
    stream_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, size_t n_tiles, Mark* marks) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * 8, w = static_cast<size_t>(blockIdx.x) * 8 + warp;
    const unsigned long long t0 = gtime();
    unsigned done = 0;
    for (size_t t = w; t < n_tiles; t += total_warps * 2) {
        uint4 a[2][3];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const size_t tt = t + u * total_warps;
#pragma unroll
            for (int j = 0; j < 3; ++j)
                a[u][j] = tt < n_tiles ? ld_stream(in + tt * 96 + j * 32 + lane) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const size_t tt = t + u * total_warps;
            if (tt >= n_tiles) break;
            uint4 r;
            r.x = a[u][0].x ^ a[u][1].y ^ a[u][2].z;
            r.y = a[u][0].y + a[u][1].z;
            r.z = a[u][0].z ^ a[u][2].w;
            r.w = a[u][0].w + a[u][1].x + a[u][2].x;
            out[tt * 32 + lane] = r;
            ++done;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) marks[blockIdx.x] = Mark{t0, gtime(), smid(), done};
}

int main(int argc, char** argv) {
    const int batches = argc > 1 ? std::atoi(argv[1]) : 20;
    const int repeats = argc > 2 ? std::atoi(argv[2]) : 5;
    const size_t n = static_cast<size_t>(batches) << 20, n_tiles = n / 32;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int grid = sms * 4;
    uint4 *in = nullptr, *out = nullptr;
    Mark* d_marks = nullptr;
    CK(cudaMalloc(&in, n * 48));
    CK(cudaMalloc(&out, n * 16));
    CK(cudaMalloc(&d_marks, grid * sizeof(Mark)));
    CK(cudaMemset(in, 1, n * 48));
    std::vector<Mark> marks(grid);
    for (int rep = 0; rep < repeats; ++rep) {
        stream_kernel<<<grid, 256>>>(in, out, n_tiles, d_marks);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(marks.data(), d_marks, grid * sizeof(Mark), cudaMemcpyDeviceToHost));
        unsigned long long first = ~0ull, last = 0, first_end = ~0ull;
        for (const Mark& m : marks) {
            first = std::min(first, m.start);
            last = std::max(last, m.end);
            first_end = std::min(first_end, m.end);
        }
        std::vector<double> per_sm(sms, 0.0);
        for (const Mark& m : marks) per_sm[m.sm % sms] = std::max(per_sm[m.sm % sms], double(m.end - first) / 1e3);
        std::vector<double> sorted(per_sm);
        std::sort(sorted.begin(), sorted.end());
        std::printf("rep %d: kernel %.1f us (%.2f TB/s), first CTA done at %.1f us, median SM %.1f us, last %.1f us "
                    "-> spread %.1f us = %.1f %%\n",
                    rep, double(last - first) / 1e3, double(n) * 64 / double(last - first) / 1e3,
                    double(first_end - first) / 1e3, sorted[sms / 2], sorted[sms - 1], double(last - first_end) / 1e3,
                    100.0 * double(last - first_end) / double(last - first));
        if (rep == repeats - 1) {
            std::printf("finish time by SM id (us), 16 per row:\n");
            for (int s = 0; s < sms; ++s) std::printf("%7.1f%s", per_sm[s], (s % 16 == 15 || s == sms - 1) ? "\n" : "");
        }
    }
    return 0;
}
