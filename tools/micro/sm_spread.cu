// Microbenchmark: do all SMs get the same share of HBM bandwidth?  148 persistent CTAs (one per SM) each stream the
// SAME number of bytes into shared memory with 1-D bulk async copies (the issue pattern that reaches the copy peak in
// bulk_copy_rate.cu) and record their SM id and their own start / end times (%globaltimer).  Prints the per-SM times
// in SM-id order for three runs: a spread that repeats SM for SM is a property of the chip (SMs share their GPC's
// port to the memory system, and GPCs do not all hold the same number of SMs), not noise.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sm_spread sm_spread.cu ; run on a B200.
#include <algorithm>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

constexpr int kWarps = 8, kLanes = 4, kSlots = 3, kCopy = 1536;      // 148 KB of ring: one CTA per SM

__global__ void __launch_bounds__(kWarps * 32, 1) k(const uint8_t* src, size_t per_cta, unsigned long long* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint8_t* ring = smem + 4096;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < kWarps * kSlots; ++i)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(&bars[i])), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const unsigned long long t0 = gtime();
    const size_t per_warp = (per_cta / kWarps) & ~(size_t)1023;
    const uint8_t* base = src + (size_t)blockIdx.x * per_cta + (size_t)warp * per_warp;
    const size_t slot_bytes = (size_t)kLanes * kCopy;
    const int n = (int)(per_warp / slot_bytes);
    uint8_t* myring = ring + (size_t)warp * kSlots * slot_bytes;
    uint64_t* mybars = bars + warp * kSlots;
    uint32_t phase = 0;
    int slot = 0;
    for (int i = 0; i < n + kSlots; ++i) {
        if (i >= kSlots)
            asm volatile("{\n\t.reg .pred p;\n\tW_%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                         ::"r"(saddr(&mybars[slot])), "r"(phase) : "memory");
        if (i < n) {
            if (lane == 0)
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(&mybars[slot])), "r"((uint32_t)slot_bytes) : "memory");
            __syncwarp();
            if (lane < kLanes)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(saddr(myring + (size_t)slot * slot_bytes + (size_t)lane * kCopy)),
                               "l"(base + (size_t)i * slot_bytes + (size_t)lane * kCopy), "r"(kCopy), "r"(saddr(&mybars[slot])) : "memory");
        }
        if (++slot == kSlots) { slot = 0; if (i >= kSlots) phase ^= 1u; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        out[3 * blockIdx.x] = smid;
        out[3 * blockIdx.x + 1] = t0;
        out[3 * blockIdx.x + 2] = gtime();
    }
}

int main() {
    const int grid = 148;
    const size_t per_cta = (size_t)8 << 20;                   // 8 MiB per SM, 1.2 GB per launch: ~190 us at the copy peak
    uint8_t* src;
    cudaMalloc(&src, per_cta * grid);
    cudaMemset(src, 1, per_cta * grid);
    unsigned long long* out;
    cudaMalloc(&out, 3 * grid * sizeof(unsigned long long));
    const size_t smem = 4096 + (size_t)kWarps * kSlots * kLanes * kCopy;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    std::vector<std::vector<double>> runs;
    for (int rep = 0; rep < 4; ++rep) {
        k<<<grid, kWarps * 32, smem>>>(src, per_cta, out);
        cudaDeviceSynchronize();
        if (cudaGetLastError() != cudaSuccess) { printf("launch failed\n"); return 1; }
        std::vector<unsigned long long> h(3 * grid);
        cudaMemcpy(h.data(), out, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        if (rep == 0) continue;                                       // warm-up
        std::vector<double> us(256, 0.0);
        unsigned long long first = ~0ull, last = 0;
        for (int b = 0; b < grid; ++b) {
            us[h[3 * b]] = (h[3 * b + 2] - h[3 * b + 1]) / 1e3;
            first = std::min(first, h[3 * b + 1]); last = std::max(last, h[3 * b + 2]);
        }
        runs.push_back(us);
        std::vector<double> v;
        for (int s = 0; s < 256; ++s) if (us[s] > 0) v.push_back(us[s]);
        std::sort(v.begin(), v.end());
        printf("run %d: %zu SMs, per-SM time min %.1f  p25 %.1f  median %.1f  p75 %.1f  max %.1f us; launch %.1f us = %.0f GB/s; if all SMs ran at the mean rate: %.1f us\n",
               rep, v.size(), v.front(), v[v.size() / 4], v[v.size() / 2], v[3 * v.size() / 4], v.back(), (last - first) / 1e3,
               per_cta * grid / ((last - first) * 1e-9) / 1e9, [&] { double r = 0; for (double x : v) r += 1.0 / x; return v.size() / r; }());
    }
    (void)smem;
    printf("per-SM time (us) by SM id, three runs:\n");
    for (int s = 0; s < 256; ++s) {
        if (runs[0][s] <= 0) continue;
        printf("sm %3d  %7.1f %7.1f %7.1f\n", s, runs[0][s], runs[1][s], runs[2][s]);
    }
    return 0;
}
