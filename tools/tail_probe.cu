// tail_probe.cu — what does a second launch of short CTAs behind a persistent grid actually do?
//
// Same streaming shape as spread_probe.cu (three coalesced 128-bit loads and one 128-bit store per
// lane and 32-record tile).  The main grid (4 CTAs of 256 threads per SM) takes the first
// (100 - tail_percent) % of the tiles in equal shares; a tail grid of short CTAs
// (tiles_per_warp tiles per warp) takes the rest, launched right behind it with programmatic
// stream serialization; the main kernel triggers its dependents at once, tail CTAs never wait.
// Printed: when the main CTAs end, when the tail CTAs start and end, how long a tail CTA lives.
//
//   ./tail_probe [batches=20] [tail_percent=8] [tiles_per_warp=4] [repeats=3]
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
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
    stream_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, size_t first_tile, size_t end_tile,
                  Mark* marks, bool wait_at_end = false) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t total_warps = static_cast<size_t>(gridDim.x) * 8, w = static_cast<size_t>(blockIdx.x) * 8 + warp;
    const unsigned long long t0 = gtime();
    unsigned done = 0;
    uint4 a[3];
    size_t t = first_tile + w;
    if (t < end_tile)
        for (int j = 0; j < 3; ++j) a[j] = ld_stream(in + t * 96 + j * 32 + lane);
    while (t < end_tile) {
        uint4 r;
        r.x = a[0].x ^ a[1].y ^ a[2].z;
        r.y = a[0].y + a[1].z;
        r.z = a[0].z ^ a[2].w;
        r.w = a[0].w + a[1].x + a[2].x;
        const size_t cur = t;
        t += total_warps;
        if (t < end_tile)
            for (int j = 0; j < 3; ++j) a[j] = ld_stream(in + t * 96 + j * 32 + lane);
        out[cur * 32 + lane] = r;
        ++done;
    }
    __syncthreads();
    if (threadIdx.x == 0) marks[blockIdx.x] = Mark{t0, gtime(), smid(), done};
    if (wait_at_end) asm volatile("griddepcontrol.wait;" ::: "memory");
}

// 55 KiB of dynamic shared memory per CTA: four CTAs fill an SM, as the dispatch kernel's registers do,
// so a tail CTA can start only where a main CTA has left.
constexpr int kPadSmem = 55 * 1024;

// ./tail_probe chain <batches per launch> <launches>: the main grid alone, `launches` times back to back with
// programmatic stream serialization; prints when each launch's first CTA starts and its last CTA ends.
int chain(int batches, int launches) {
    const size_t n = static_cast<size_t>(batches) << 20, n_tiles = n / 32;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int grid = sms * 4;
    uint4 *in = nullptr, *out = nullptr;
    Mark* d_marks = nullptr;
    CK(cudaMalloc(&in, n * 48));
    CK(cudaMalloc(&out, n * 16));
    CK(cudaMalloc(&d_marks, static_cast<size_t>(launches) * grid * sizeof(Mark)));
    CK(cudaMemset(in, 1, n * 48));
    cudaStream_t stream;
    CK(cudaStreamCreate(&stream));
    for (int rep = 0; rep < 2; ++rep) {
        for (int l = 0; l < launches; ++l) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(grid);
            cfg.blockDim = dim3(256);
            cfg.stream = stream;
            cfg.dynamicSmemBytes = kPadSmem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = l ? 1 : 0;
            CK(cudaLaunchKernelEx(&cfg, stream_kernel, static_cast<const uint4*>(in), out, size_t(0), n_tiles,
                                  d_marks + static_cast<size_t>(l) * grid, true));
        }
        CK(cudaStreamSynchronize(stream));
    }
    std::vector<Mark> marks(static_cast<size_t>(launches) * grid);
    CK(cudaMemcpy(marks.data(), d_marks, marks.size() * sizeof(Mark), cudaMemcpyDeviceToHost));
    unsigned long long t0 = ~0ull, prev_end = 0;
    for (int i = 0; i < grid; ++i) t0 = std::min(t0, marks[i].start);
    for (int l = 0; l < launches; ++l) {
        unsigned long long first = ~0ull, first_end = ~0ull, last_start = 0, last = 0;
        for (int i = 0; i < grid; ++i) {
            const Mark& m = marks[static_cast<size_t>(l) * grid + i];
            first = std::min(first, m.start);
            last_start = std::max(last_start, m.start);
            first_end = std::min(first_end, m.end);
            last = std::max(last, m.end);
        }
        std::printf("launch %d: CTAs start %.1f .. %.1f us, end %.1f .. %.1f us; since the previous launch ended: %.1f us "
                    "= %.2f us per batch\n", l, double(first - t0) / 1e3, double(last_start - t0) / 1e3,
                    double(first_end - t0) / 1e3, double(last - t0) / 1e3, double(last - (l ? prev_end : t0)) / 1e3,
                    double(last - (l ? prev_end : t0)) / 1e3 / batches);
        prev_end = last;
    }
    return 0;
}

int main(int argc, char** argv) {
    CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPadSmem));
    if (argc > 3 && std::string(argv[1]) == "chain") return chain(std::atoi(argv[2]), std::atoi(argv[3]));
    const int batches = argc > 1 ? std::atoi(argv[1]) : 20;
    const int tail_percent = argc > 2 ? std::atoi(argv[2]) : 8;
    const int tiles_per_warp = argc > 3 ? std::atoi(argv[3]) : 4;
    const int repeats = argc > 4 ? std::atoi(argv[4]) : 3;
    const size_t n = static_cast<size_t>(batches) << 20, n_tiles = n / 32;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int grid = sms * 4;
    const size_t tail_tiles = n_tiles * tail_percent / 100, main_tiles = n_tiles - tail_tiles;
    const int tail_grid = tail_tiles ? static_cast<int>((tail_tiles + 8 * tiles_per_warp - 1) / (8 * tiles_per_warp)) : 0;
    uint4 *in = nullptr, *out = nullptr;
    Mark *d_main = nullptr, *d_tail = nullptr;
    CK(cudaMalloc(&in, n * 48));
    CK(cudaMalloc(&out, n * 16));
    CK(cudaMalloc(&d_main, grid * sizeof(Mark)));
    CK(cudaMalloc(&d_tail, (tail_grid + 1) * sizeof(Mark)));
    CK(cudaMemset(in, 1, n * 48));
    cudaStream_t stream;
    CK(cudaStreamCreate(&stream));
    std::vector<Mark> m_main(grid), m_tail(tail_grid);
    for (int rep = 0; rep < repeats; ++rep) {
        stream_kernel<<<grid, 256, kPadSmem, stream>>>(in, out, 0, main_tiles, d_main, false);
        if (tail_grid) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(tail_grid);
            cfg.blockDim = dim3(256);
            cfg.stream = stream;
            cfg.dynamicSmemBytes = kPadSmem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            CK(cudaLaunchKernelEx(&cfg, stream_kernel, static_cast<const uint4*>(in), out, main_tiles, n_tiles, d_tail, false));
        }
        CK(cudaStreamSynchronize(stream));
        CK(cudaMemcpy(m_main.data(), d_main, grid * sizeof(Mark), cudaMemcpyDeviceToHost));
        if (tail_grid) CK(cudaMemcpy(m_tail.data(), d_tail, tail_grid * sizeof(Mark), cudaMemcpyDeviceToHost));
        unsigned long long first = ~0ull, main_first_end = ~0ull, main_last_end = 0, tail_first_start = ~0ull,
                           tail_last_end = 0;
        for (const Mark& m : m_main) {
            first = std::min(first, m.start);
            main_first_end = std::min(main_first_end, m.end);
            main_last_end = std::max(main_last_end, m.end);
        }
        double life = 0;
        std::vector<int> per_sm(sms, 0);
        for (const Mark& m : m_tail) {
            tail_first_start = std::min(tail_first_start, m.start);
            tail_last_end = std::max(tail_last_end, m.end);
            life += double(m.end - m.start);
            ++per_sm[m.sm % sms];
        }
        const unsigned long long last = std::max(main_last_end, tail_last_end);
        std::printf("rep %d: all done at %.1f us (%.2f TB/s); main CTAs end %.1f .. %.1f us", rep, double(last - first) / 1e3,
                    double(n) * 64 / double(last - first) / 1e3, double(main_first_end - first) / 1e3,
                    double(main_last_end - first) / 1e3);
        if (tail_grid) {
            std::sort(per_sm.begin(), per_sm.end());
            std::printf("; %d tail CTAs: first starts %.1f us, last ends %.1f us, mean life %.2f us, per SM %d .. %d (median %d)",
                        tail_grid, double(tail_first_start - first) / 1e3, double(tail_last_end - first) / 1e3,
                        life / tail_grid / 1e3, per_sm.front(), per_sm.back(), per_sm[sms / 2]);
        }
        std::printf("\n");
    }
    return 0;
}
