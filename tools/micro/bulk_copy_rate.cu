// Microbenchmark: how fast can ONE SM pull global memory into shared memory with 1-D bulk async copies
// (cp.async.bulk, SASS UBLKCP), as a function of copy size, copies in flight and the number of issuing warps?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_copy_rate bulk_copy_rate.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(1024, 1) k(const uint8_t* src, size_t per_cta, int copy_bytes, int slots, int warps, int lanes) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);           // [warps][slots]
    uint8_t* ring = smem + 4096;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < warps * slots; ++i)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(&bars[i])), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp >= warps) return;
    // this warp's share of the CTA's bytes, in groups of `lanes` copies per slot
    const size_t per_warp = (per_cta / warps) & ~(size_t)1023;
    const uint8_t* base = src + (size_t)blockIdx.x * per_cta + (size_t)warp * per_warp;
    const size_t slot_bytes = (size_t)lanes * copy_bytes;
    const int n = (int)(per_warp / slot_bytes);
    uint8_t* myring = ring + (size_t)warp * slots * slot_bytes;
    uint64_t* mybars = bars + warp * slots;
    uint32_t phase = 0;
    int slot = 0;
    for (int i = 0; i < n + slots; ++i) {
        if (i >= slots) {       // wait for the copy issued `slots` steps ago
            asm volatile("{\n\t.reg .pred p;\n\tW_%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                         ::"r"(saddr(&mybars[slot])), "r"(phase) : "memory");
        }
        if (i < n) {
            if (lane == 0)
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(&mybars[slot])), "r"((uint32_t)slot_bytes) : "memory");
            __syncwarp();
            if (lane < lanes)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(saddr(myring + (size_t)slot * slot_bytes + (size_t)lane * copy_bytes)),
                               "l"(base + (size_t)i * slot_bytes + (size_t)lane * copy_bytes), "r"(copy_bytes), "r"(saddr(&mybars[slot])) : "memory");
        }
        if (++slot == slots) { slot = 0; if (i >= slots) phase ^= 1u; }
    }
}

int main() {
    const size_t total = (size_t)2 << 30;
    uint8_t* src;
    cudaMalloc(&src, total);
    cudaMemset(src, 1, total);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int grid = 148;
    const size_t per_cta = (total / grid) & ~(size_t)((1 << 20) - 1);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("copy_bytes lanes warps slots  inflight_KB  GB/s\n");
    const int cfgs[][4] = {
        // copy_bytes, lanes per slot, warps, slots
        {1536, 4, 1, 4}, {1536, 4, 1, 8}, {1536, 4, 1, 16}, {1536, 4, 1, 32},
        {1536, 8, 1, 8}, {1536, 16, 1, 4}, {1536, 32, 1, 2}, {1536, 32, 1, 4},
        {1536, 4, 2, 8}, {1536, 4, 4, 8}, {1536, 4, 8, 4}, {1536, 4, 16, 2},
        {688, 8, 10, 2}, {688, 8, 10, 3}, {688, 8, 20, 2},
        {6144, 1, 1, 8}, {6144, 1, 1, 16}, {6144, 1, 1, 32}, {6144, 1, 4, 8},
        {16384, 1, 1, 4}, {16384, 1, 1, 8}, {16384, 1, 1, 12}, {16384, 1, 4, 3},
        {512, 8, 8, 4}, {256, 16, 8, 4}, {128, 32, 8, 4},
    };
    for (auto& c : cfgs) {
        const int cb = c[0], lanes = c[1], warps = c[2], slots = c[3];
        const size_t smem = 4096 + (size_t)warps * slots * lanes * cb;
        if (smem > 220 * 1024) { printf("%d %d %d %d skipped (smem)\n", cb, lanes, warps, slots); continue; }
        float best = 1e9f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            k<<<grid, 32 * warps, smem>>>(src, per_cta, cb, slots, warps, lanes);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        cudaError_t err = cudaGetLastError();
        if (err != cudaSuccess) { printf("error %s\n", cudaGetErrorString(err)); return 1; }
        const size_t moved = (((per_cta / warps) & ~(size_t)1023) / ((size_t)lanes * cb)) * (size_t)lanes * cb * warps * grid;
        printf("%6d %5d %5d %5d  %8.1f  %8.1f\n", cb, lanes, warps, slots, (double)warps * slots * lanes * cb / 1024.0, moved / (best * 1e-3) / 1e9);
    }
    return 0;
}
