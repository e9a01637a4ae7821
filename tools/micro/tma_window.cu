// Micro test: does a rank-4 uint8 tensor map with an OVERLAPPING dimension (stride smaller than the extent of the
// dimension below it) fetch an exact crop window as [rows][row_bytes] with one cp.async.bulk.tensor?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_window tma_window.cu   (no -lcuda: entry point via the runtime)
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void k(const __grid_constant__ CUtensorMap map, int c0, int c2, int c3, int bytes, uint8_t* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(saddr(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(&bar)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                     ::"r"(saddr(smem)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(c0), "r"(0), "r"(c2), "r"(c3), "r"(saddr(&bar)) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred p;\n\tW_%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                 ::"r"(saddr(&bar)) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = smem[i];
}

__global__ void k2(const __grid_constant__ CUtensorMap map, int c0, int c1, int bytes, uint8_t* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(saddr(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(&bar)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(saddr(smem)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(c0), "r"(c1), "r"(saddr(&bar)) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred p;\n\tW_%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                 ::"r"(saddr(&bar)) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = smem[i];
}

template <int R>
__global__ void kr(const __grid_constant__ CUtensorMap map, int c0, int c1, int c2, int c3, int bytes, uint8_t* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(saddr(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(&bar)), "r"(bytes) : "memory");
        if (R == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(saddr(smem)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(c0), "r"(c1), "r"(c2), "r"(saddr(&bar)) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                         ::"r"(saddr(smem)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(saddr(&bar)) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred p;\n\tW_%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                 ::"r"(saddr(&bar)) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = smem[i];
}

typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                        const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// one variant: crop rows of out_w * 3 bytes cut into kk pieces; dims0 = the extent declared for dimension 0
int run(Enc enc, const char* name, int out_w, int kk, bool tight_dim0, int dim1_stride = 0, int l2 = 1) {
    const int B = 3, H = 64, W = 256, rows = 8;
    const int pitch = W * 3, row_bytes = out_w * 3, inner = row_bytes / kk;
    std::vector<uint8_t> h((size_t)B * H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size());
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    cudaMalloc(&o, rows * row_bytes);
    CUtensorMap map;
    const cuuint64_t dims[4] = {(cuuint64_t)(tight_dim0 ? pitch - (kk - 1) * inner : pitch), (cuuint64_t)kk, (cuuint64_t)H, (cuuint64_t)B};
    const cuuint64_t strides[3] = {(cuuint64_t)(dim1_stride ? dim1_stride : inner), (cuuint64_t)pitch, (cuuint64_t)H * pitch};
    const cuuint32_t box[4] = {(cuuint32_t)inner, (cuuint32_t)kk, (cuuint32_t)rows, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, l2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("%-40s encode %d", name, (int)r);
    if (r != CUDA_SUCCESS) { printf("\n"); return 1; }
    const int left = 7, top = 13, b = 2;
    k<<<1, 128, rows * row_bytes>>>(map, left * 3, top, b, rows * row_bytes, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  kernel: %s", cudaGetErrorString(e));
    if (e != cudaSuccess) { printf("\n"); return 2; }
    std::vector<uint8_t> got(rows * row_bytes);
    cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < rows; ++rr)
        for (int x = 0; x < row_bytes; ++x)
            bad += got[rr * row_bytes + x] != h[((size_t)b * H + top + rr) * pitch + left * 3 + x];
    printf("  mismatches: %d of %d\n", bad, rows * row_bytes);
    return bad != 0;
}

int run2(Enc enc, CUtensorMapDataType dt, int esize, const char* name) {
    const int H = 64, W = 256, rows = 8, pitch = W * 3, box_bytes = 192;
    std::vector<uint8_t> h((size_t)H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size());
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    cudaMalloc(&o, rows * box_bytes);
    CUtensorMap map;
    const cuuint64_t dims[2] = {(cuuint64_t)(pitch / esize), (cuuint64_t)H};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {(cuuint32_t)(box_bytes / esize), (cuuint32_t)rows};
    const cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&map, dt, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("%-40s encode %d", name, (int)r);
    if (r != CUDA_SUCCESS) { printf("\n"); return 1; }
    k2<<<1, 128, rows * box_bytes>>>(map, 48 / esize, 13, rows * box_bytes, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  kernel: %s", cudaGetErrorString(e));
    if (e != cudaSuccess) { printf("\n"); return 2; }
    std::vector<uint8_t> got(rows * box_bytes);
    cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < rows; ++rr)
        for (int x = 0; x < box_bytes; ++x) bad += got[rr * box_bytes + x] != h[(size_t)(13 + rr) * pitch + 48 + x];
    printf("  mismatches: %d of %d\n", bad, rows * box_bytes);
    return bad != 0;
}

// plain maps of a [B][H][pitch] byte array: rank 3 = (pitch, H, B); rank 4 = (256, 3, H, B) (a row as three 256-byte thirds)
int run_plain(Enc enc, int rank, const char* name) {
    const int B = 3, H = 64, pitch = 768, rows = 8, box_bytes = 192;
    std::vector<uint8_t> h((size_t)B * H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size());
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    cudaMalloc(&o, rows * box_bytes);
    CUtensorMap map;
    CUresult r;
    const cuuint32_t es[4] = {1, 1, 1, 1};
    if (rank == 3) {
        const cuuint64_t dims[3] = {768, 64, 3}, strides[2] = {768, 49152};
        const cuuint32_t box[3] = {192, 8, 1};
        r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        const cuuint64_t dims[4] = {256, 3, 64, 3}, strides[3] = {256, 768, 49152};
        const cuuint32_t box[4] = {192, 1, 8, 1};
        r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    printf("%-40s encode %d", name, (int)r);
    if (r != CUDA_SUCCESS) { printf("\n"); return 1; }
    // window: image 2, rows 13.., bytes 48.. of the (first third of the) row
    if (rank == 3) kr<3><<<1, 128, rows * box_bytes>>>(map, 48, 13, 2, 0, rows * box_bytes, o);
    else kr<4><<<1, 128, rows * box_bytes>>>(map, 48, 0, 13, 2, rows * box_bytes, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  kernel: %s", cudaGetErrorString(e));
    if (e != cudaSuccess) { printf("\n"); return 2; }
    std::vector<uint8_t> got(rows * box_bytes);
    cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < rows; ++rr)
        for (int x = 0; x < box_bytes; ++x) bad += got[rr * box_bytes + x] != h[((size_t)2 * H + 13 + rr) * pitch + 48 + x];
    printf("  mismatches: %d of %d\n", bad, rows * box_bytes);
    return bad != 0;
}

// the window map: dimension 1 steps by `inner` bytes INSIDE dimension 0's extent (overlap), box = one whole crop row per row
int run_window(Enc enc, int inner, int kk, int dim0, const char* name) {
    const int B = 3, H = 64, pitch = 768, rows = 8, row_bytes = inner * kk;
    std::vector<uint8_t> h((size_t)B * H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size());
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    cudaMalloc(&o, rows * row_bytes);
    CUtensorMap map;
    const cuuint32_t es[4] = {1, 1, 1, 1};
    const cuuint64_t dims[4] = {(cuuint64_t)dim0, (cuuint64_t)kk, 64, 3}, strides[3] = {(cuuint64_t)inner, 768, 49152};
    const cuuint32_t box[4] = {(cuuint32_t)inner, (cuuint32_t)kk, 8, 1};
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("%-40s encode %d", name, (int)r);
    if (r != CUDA_SUCCESS) { printf("\n"); return 1; }
    const int left3 = 21;
    kr<4><<<1, 128, rows * row_bytes>>>(map, left3, 0, 13, 2, rows * row_bytes, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  kernel: %s", cudaGetErrorString(e));
    if (e != cudaSuccess) { printf("\n"); return 2; }
    std::vector<uint8_t> got(rows * row_bytes);
    cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < rows; ++rr)
        for (int x = 0; x < row_bytes; ++x) bad += got[rr * row_bytes + x] != h[((size_t)2 * H + 13 + rr) * pitch + left3 + x];
    printf("  mismatches: %d of %d\n", bad, rows * row_bytes);
    return bad != 0;
}

int main(int argc, char** argv) {
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    const int which = argc > 1 ? atoi(argv[1]) : 0;       // one variant per process: a fault poisons the context
    if (which == 0) return run((Enc)fp, "64 px rows, 1 piece (no overlap)", 64, 1, true);
    if (which == 1) return run((Enc)fp, "64 px rows, 1 piece, dim0 = pitch", 64, 1, false);
    if (which == 2) return run((Enc)fp, "224 px rows, 3 pieces, dim0 tight", 224, 3, true);
    if (which == 3) return run((Enc)fp, "224 px rows, 3 pieces, dim0 = pitch", 224, 3, false);
    if (which == 6) return run((Enc)fp, "rank 4, 1 piece, dim1 stride = pitch (no overlap)", 64, 1, false, 768);
    if (which == 7) return run((Enc)fp, "rank 4, 1 piece, overlap, L2 promotion none", 64, 1, true, 0, 0);
    if (which == 10) return run_window((Enc)fp, 224, 3, 768 - 2 * 224, "window 3 x 224, dim0 tight (320)");
    if (which == 11) return run_window((Enc)fp, 224, 3, 768, "window 3 x 224, dim0 = pitch");
    if (which == 12) return run_window((Enc)fp, 192, 1, 768, "window 1 x 192, dim0 = pitch");
    if (which == 13) return run_window((Enc)fp, 224, 3, 256, "window 3 x 224, dim0 = 256");
    if (which == 8) return run_plain((Enc)fp, 3, "rank 3 plain (pitch, H, B)");
    if (which == 9) return run_plain((Enc)fp, 4, "rank 4 plain (256, 3, H, B)");
    if (which == 4) return run2((Enc)fp, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, "rank 2, uint8, box 192 x 8");
    if (which == 5) return run2((Enc)fp, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, "rank 2, uint32, box 48 x 8");
    return 0;
}
