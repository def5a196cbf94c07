// update.cu — in-place registry updates on the device.
//
// One node joining or leaving a service (one etcd key appearing or expiring,
// cluster/registry.go:58-83,119-150) changes exactly two things in the image: the service's
// node count, and - because global node ids are ranks in the sorted key range - the node_base
// of every service that sorts after it.  Rebuilding and re-uploading the image for that costs
// milliseconds of host work for a 10 000-service registry; patching a copy of the slot table on
// the device costs microseconds and needs no host data at all.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ptype_b200.h"
#include "device_image.h"
#include "table_hash.h"

namespace ptype {
namespace {

// One CTA.  Phase 1: find the service's slot among its three cuckoo candidates.  Phase 2: walk
// the table: the service's own count changes by `delta`, every service that sorts after it in
// etcd key order (order[]: the rank the image builder recorded per slot) moves its base by
// `delta`.  Ranks, not bases, decide: node-less services share their base with a neighbour.
// result[0] = 0 ok, 1 unknown service, 2 count would leave [0, 2^24).
__global__ void __launch_bounds__(1024)
    patch_service_kernel(ptype_slot* __restrict__ slots, const uint32_t* __restrict__ order,
                         const uint32_t n_slots, const uint32_t seed,
                         const uint32_t key_lo, const uint32_t key_hi, const int delta,
                         uint32_t* __restrict__ result) {
    __shared__ uint32_t s_at, s_n;
    if (threadIdx.x == 0) s_at = PTYPE_SLOT_EMPTY;
    __syncthreads();
    if (threadIdx.x < PTYPE_SLOT_WAYS) {
        const uint32_t at = slot_of(key_lo, key_hi, seed, static_cast<int>(threadIdx.x), n_slots);
        const ptype_slot s = slots[at];
        if (s.node_base != PTYPE_SLOT_EMPTY && static_cast<uint32_t>(s.key) == key_lo &&
            static_cast<uint32_t>(s.key >> 32) == key_hi) {
            s_at = at;  // duplicates among the candidates name the same slot
            s_n = s.n_and_recv & 0xFFFFFFu;
        }
    }
    __syncthreads();
    if (s_at == PTYPE_SLOT_EMPTY) {
        if (threadIdx.x == 0) result[0] = 1;
        return;
    }
    const long long n_new = static_cast<long long>(s_n) + delta;
    if (n_new < 0 || n_new > 0xFFFFFF) {
        if (threadIdx.x == 0) result[0] = 2;
        return;
    }
    const uint32_t rank = order[s_at];
    for (uint32_t i = threadIdx.x; i < n_slots; i += blockDim.x) {
        ptype_slot s = slots[i];
        if (s.node_base == PTYPE_SLOT_EMPTY) continue;
        if (i == s_at) {
            s.n_and_recv = (s.n_and_recv & 0xFF000000u) | static_cast<uint32_t>(n_new);
            slots[i] = s;
        } else if (order[i] > rank) {
            s.node_base = static_cast<uint32_t>(static_cast<long long>(s.node_base) + delta);
            slots[i] = s;
        }
    }
    if (threadIdx.x == 0) result[0] = 0;
}

}  // namespace

cudaError_t launch_patch_service(ptype_slot* slots, const uint32_t* order, uint32_t n_slots, uint32_t seed,
                                 uint64_t key, int delta, uint32_t* result, cudaStream_t stream) {
    patch_service_kernel<<<1, 1024, 0, stream>>>(slots, order, n_slots, seed, static_cast<uint32_t>(key),
                                                  static_cast<uint32_t>(key >> 32), delta, result);
    return cudaGetLastError();
}

}  // namespace ptype
