// partial_write_probe.cu — what does HBM traffic look like when a kernel writes 8 bytes into every
// 48-byte record of a large array (ptype_req.seq, written by assign_seq) without reading the rest?
// Run under `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`: the read column tells whether
// partially written sectors are filled from DRAM before they are written back.
//   ./partial_write_probe [records=16777216]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__global__ void write_seq(unsigned long long* recs, size_t n) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        recs[i * 6 + 1] = i;  // 8 bytes at offset 8 of a 48-byte record
}
__global__ void write_sector(uint4* recs, size_t n) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        recs[i * 3] = make_uint4(1, 2, static_cast<unsigned>(i), 0);      // bytes 0..15
        recs[i * 3 + 1] = make_uint4(3, 4, 5, 6);                         // bytes 16..31: first sector complete
    }
}
__global__ void read_key(const unsigned long long* recs, size_t n, unsigned long long* out) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    unsigned long long acc = 0;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) acc += recs[i * 6];
    if (acc == 12345) *out = acc;
}

int main(int argc, char** argv) {
    const size_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : (1u << 24);
    void* buf = nullptr;
    unsigned long long* out = nullptr;
    if (cudaMalloc(&buf, n * 48) != cudaSuccess || cudaMalloc(&out, 8) != cudaSuccess) return 1;
    cudaMemset(buf, 0, n * 48);
    cudaDeviceSynchronize();
    cudaEvent_t e[4];
    for (auto& x : e) cudaEventCreate(&x);
    cudaEventRecord(e[0]);
    write_seq<<<148 * 8, 256>>>(static_cast<unsigned long long*>(buf), n);
    cudaEventRecord(e[1]);
    write_sector<<<148 * 8, 256>>>(static_cast<uint4*>(buf), n);
    cudaEventRecord(e[2]);
    read_key<<<148 * 8, 256>>>(static_cast<unsigned long long*>(buf), n, out);
    cudaEventRecord(e[3]);
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    float a, b, c;
    cudaEventElapsedTime(&a, e[0], e[1]);
    cudaEventElapsedTime(&b, e[1], e[2]);
    cudaEventElapsedTime(&c, e[2], e[3]);
    std::printf("records %zu: write 8 B/record %.1f us, write 32 B/record %.1f us, read 8 B/record %.1f us\n", n, a * 1e3,
                b * 1e3, c * 1e3);
    return 0;
}
