// Shared helpers of the C-ABI library: error reporting, launch accounting, small PTX wrappers.
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/torchbooster_b200.h"

// Thread-local message for the last non-zero return (tb_abi.cu).
int tb_fail(int code, const char* fmt, ...);
int tb_cuda_fail(cudaError_t err, const char* what);
void tb_count_launch(int n = 1);

#define TB_REQUIRE(cond, ...)                          \
    do {                                               \
        if (!(cond)) return tb_fail(TB_EINVAL, __VA_ARGS__); \
    } while (0)

#define TB_CUDA(expr)                                  \
    do {                                               \
        cudaError_t _e = (expr);                       \
        if (_e != cudaSuccess) return tb_cuda_fail(_e, #expr); \
    } while (0)

// Peek (do not clear sticky state of other libraries) after a launch.
#define TB_LAUNCH_CHECK(name)                          \
    do {                                               \
        cudaError_t _e = cudaGetLastError();           \
        if (_e != cudaSuccess) return tb_cuda_fail(_e, name); \
        tb_count_launch();                             \
    } while (0)

static inline int tb_dtype_size(int dt) { return dt == TB_F32 ? 4 : ((dt == TB_BF16 || dt == TB_F16) ? 2 : 0); }

struct TbDevice {
    int device;
    int sm_count;
    int max_smem_optin;
};
// Cached per-device attributes (read-only after first use).
int tb_device(TbDevice* out);

// ---------------------------------------------------------------------------------------------
// device-side helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t tb_smem_addr(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void tb_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tb_smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void tb_mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tb_mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tb_smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void tb_mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tb_smem_addr(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tb_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(tb_smem_addr(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine; SASS UBLKCP), completion on an mbarrier.
// src and dst 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tb_bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     tb_smem_addr(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(tb_smem_addr(bar))
                 : "memory");
}

// Generic-proxy reads of a shared-memory stage are ordered before the async-proxy (bulk copy)
// writes that refill it.
__device__ __forceinline__ void tb_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Programmatic dependent launch.  `wait` blocks until every kernel this launch depends on has
// completed and its writes are visible (a no-op for a normal launch); it must precede the first
// global-memory access.  `launch_dependents` lets the next PDL launch in the stream begin its
// own pre-wait preamble.
__device__ __forceinline__ void tb_grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void tb_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void tb_st_cs_v4(void* p, uint4 v) {
    asm volatile("st.global.cs.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
#endif  // __CUDACC__
