// Image-resident augmentation chain (sm_100a): pad -> crop -> flip -> [per-sample operation list]
// -> table lookup -> cast, uint8 NHWC -> NCHW, ONE launch per batch.
//
// Replaces the middle of the reference's CIFAR pipeline
// (examples/img_cls/resnet/resnet.py:96-103: RandomCrop, RandomHorizontalFlip, RandomRotation(15),
// RandAugment(), ToTensor, Normalize), which torchvision runs per sample on PIL images: every
// stage there is a full pass over the image through Pillow's C code plus Python object churn.
// Here one CTA owns one image.  The cropped / flipped image is gathered from global memory ONCE
// into shared memory (planar, the layout of the output), every operation of that sample's list
// runs out of shared memory (two buffers, ping-pong for the operations that move pixels), and
// the result leaves through the per-channel table as coalesced planar stores.  Global traffic is
// the algorithmic minimum: source bytes in, output elements out.  An image too large for one SM
// (ImageNet-sized crops) is owned by a thread-block CLUSTER of 2, 4 or 8 CTAs instead: the two
// buffers are cut into equal chunks across the cluster's shared memories and every thread reaches
// any pixel through distributed shared memory; histograms and sums are kept per CTA and added up
// by each CTA from its peers.
//
// The arithmetic of each operation is Pillow 12's, bit for bit (oracle/augment_oracle.py):
//   AFFINE        nearest neighbour in 16.16 fixed point (Geometry.c affine_fixed): rotations,
//                 shears, integer translations; outside = 0
//   BRIGHTNESS / COLOR / CONTRAST / SHARPNESS
//                 ImagingBlend(degenerate, image, alpha) = deg + alpha * (image - deg) in float32
//                 (two roundings, no FMA), truncated, clipped; degenerate = black / grey value
//                 (19595 R + 38470 G + 7471 B + 0x8000) >> 16 / grey mean / 3x3 SMOOTH
//   POSTERIZE, SOLARIZE, INVERT   byte maps
//   AUTOCONTRAST, EQUALIZE        per-band histogram -> table (ImageOps), float64 where Python is
#include <cooperative_groups.h>

#include "tb_common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int kAugThreads = 256;      // threads per CTA in a full grid; 512 / 1024 when there are few images (thread t < 256 owns histogram bin t)
constexpr int kAugMaxThreads = 1024;
constexpr int kAugChunk = 24 * 1024;

struct AugParams {
    const uint8_t* src;
    int B, H, W, C;
    int64_t bs, pitch;
    void* dst;
    int out_h, out_w;
    const int32_t* top;
    const int32_t* left;
    const uint8_t* flip;
    int pad_y, pad_x, pad_mode;
    const int32_t* ops;
    int n_ops;
    const void* lut;
    int64_t px_stride, ch_stride;   // source bytes between pixels of a row / between channels of a pixel
    int plane;          // out_h * out_w
    int buf_bytes;      // bytes of one image buffer held by one CTA (multiple of 16): the whole image, or its chunk
    uint64_t magic;     // ceil(2^40 / buf_bytes) (cluster launches)
    uint64_t wmagic;    // ceil(2^40 / out_w): pixel index -> row without a division
};

__device__ __forceinline__ int aug_fold(int i, int n, int mode, bool& inside) {
    if ((unsigned)i < (unsigned)n) return i;
    if (mode == TB_PAD_REFLECT)
        i = i < 0 ? -i : 2 * (n - 1) - i;
    else if (mode == TB_PAD_SYMMETRIC)
        i = i < 0 ? -i - 1 : 2 * n - 1 - i;
    else if (mode == TB_PAD_EDGE)
        i = i < 0 ? 0 : n - 1;
    else
        i = -1;
    if ((unsigned)i >= (unsigned)n) {
        inside = false;
        return 0;
    }
    return i;
}

// ImagingBlend: float32 multiply, float32 add (separately rounded), truncation; the clip is a
// no-op for alpha in [0, 1], where Pillow does not clip.
__device__ __forceinline__ uint8_t aug_blend(int deg, int v, float alpha) {
    const float t = __fadd_rn((float)deg, __fmul_rn(alpha, (float)(v - deg)));
    return t <= 0.0f ? 0 : (t >= 255.0f ? 255 : (uint8_t)(int)t);
}

// channel of planar element i (C <= 3; elements past the image count as the last channel)
__device__ __forceinline__ int aug_channel(int i, int plane) { return (i >= plane) + (i >= 2 * plane); }

__device__ __forceinline__ int aug_grey(int r, int g, int b) { return (r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16; }

// A planar image buffer in shared memory: whole in this CTA (CL = false), or -- for images that do
// not fit in one SM's shared memory twice -- cut into equal chunks over the CTAs of a thread-block
// cluster and reached through distributed shared memory (CL = true): element i lives in CTA
// i / chunk of the cluster at the same local offset in every CTA.
template <bool CL>
struct AugBuf;
template <>
struct AugBuf<false> {
    uint8_t* base;
    __device__ __forceinline__ int ld(int i) const { return base[i]; }
    __device__ __forceinline__ void st(int i, uint8_t v) const { base[i] = v; }
    __device__ __forceinline__ uint32_t* word(int i) const { return (uint32_t*)(base + i); }     // i % 4 == 0
};
template <>
struct AugBuf<true> {
    uint8_t* base;
    uint32_t chunk;
    uint64_t magic;     // ceil(2^40 / chunk): i / chunk == (i * magic) >> 40 for every i < 2^20
    __device__ __forceinline__ uint8_t* at(int i) const {
        const uint32_t r = (uint32_t)(((uint64_t)(uint32_t)i * magic) >> 40);
        return cg::this_cluster().map_shared_rank(base, r) + ((uint32_t)i - r * chunk);
    }
    __device__ __forceinline__ int ld(int i) const { return *at(i); }
    __device__ __forceinline__ void st(int i, uint8_t v) const { *at(i) = v; }
    __device__ __forceinline__ uint32_t* word(int i) const { return (uint32_t*)at(i); }     // chunks are multiples of 16
};

// Byte-wise map over a buffer, four bytes per access (the padding bytes after the image ride along).
template <bool CL, typename F>
__device__ __forceinline__ void aug_map(const AugBuf<CL>& buf, int total, int gt, int gn, F f) {
    for (int i = gt * 4; i < total; i += gn * 4) {
        uint32_t* w = buf.word(i);
        const uint32_t v = *w;
        *w = (uint32_t)f(i, (int)(v & 0xffu)) | ((uint32_t)f(i + 1, (int)((v >> 8) & 0xffu)) << 8) |
             ((uint32_t)f(i + 2, (int)((v >> 16) & 0xffu)) << 16) | ((uint32_t)f(i + 3, (int)(v >> 24)) << 24);
    }
}

template <bool CL>
__device__ __forceinline__ void aug_sync() {
    if constexpr (CL) cg::this_cluster().sync();
    else __syncthreads();
}
// value of a per-CTA shared variable in CTA r of the cluster
template <bool CL>
__device__ __forceinline__ int aug_peer(int* var, int r) {
    if constexpr (CL) return *cg::this_cluster().map_shared_rank(var, r);
    else return *var;
}

template <typename T, bool CL>
__global__ void __launch_bounds__(kAugMaxThreads) tb_augment_kernel(AugParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int s_hist[3 * 256];     // this CTA's share of the histogram
    __shared__ int s_full[3 * 256];     // the image's histogram
    __shared__ uint8_t s_tab[3 * 256];
    __shared__ int s_lo[3], s_hi[3], s_sum, s_warp[8];
    __shared__ int32_t s_ops[TB_AUG_MAX_OPS * 8];

    const int S = CL ? (int)cg::this_cluster().num_blocks() : 1;        // CTAs that share the image
    const int rank = CL ? (int)cg::this_cluster().block_rank() : 0;
    AugBuf<CL> cur, alt;
    cur.base = smem;
    alt.base = smem + p.buf_bytes;
    if constexpr (CL) {
        cur.chunk = alt.chunk = (uint32_t)p.buf_bytes;
        cur.magic = alt.magic = p.magic;
    }
    T* s_lut = (T*)(smem + 2 * p.buf_bytes);

    tb_grid_dependency_wait();
    const int tid = threadIdx.x;
    const int b = blockIdx.x / S;
    const int nt = (int)blockDim.x;
    const int gt = rank * nt + tid, gn = S * nt;      // this thread among the image's threads
    const int C = p.C, plane = p.plane, W = p.out_w, H = p.out_h;
    const int total = C * plane;
    const uint64_t wmagic = p.wmagic;

    {
        const T* __restrict__ glut = (const T*)p.lut;
        for (int i = tid; i < C * 256; i += nt) s_lut[i] = glut[i];
        if (tid < p.n_ops * 8) s_ops[tid] = p.ops[(int64_t)b * p.n_ops * 8 + tid];      // one round trip for the whole list
    }
    if constexpr (CL) cg::this_cluster().sync();        // every CTA of the cluster is resident before its memory is written
    // ---- gather: pad / crop / flip, interleaved source -> planar shared --------------------
    {
        const int top = (p.top ? p.top[b] : 0) - p.pad_y;
        const int left = (p.left ? p.left[b] : 0) - p.pad_x;
        const bool flip = p.flip ? (p.flip[b] != 0) : false;
        const uint8_t* __restrict__ img = p.src + (int64_t)b * p.bs;
        constexpr int GU = 4;            // pixels whose loads are in flight together
        for (int q0 = gt; q0 < plane; q0 += GU * gn) {
            uint8_t px[GU][3];
#pragma unroll
            for (int k = 0; k < GU; ++k) {
                const int q = min(q0 + k * gn, plane - 1);
                const int y = (int)(((uint64_t)(uint32_t)q * wmagic) >> 40), x = q - y * W;      // q / W, exact for q < 2^20
                const int xs = flip ? W - 1 - x : x;
                bool inside = true;
                const int u = aug_fold(top + y, p.H, p.pad_mode, inside);
                const int v = aug_fold(left + xs, p.W, p.pad_mode, inside);
                const uint8_t* __restrict__ s = img + (int64_t)u * p.pitch + (int64_t)v * p.px_stride;
#pragma unroll
                for (int c = 0; c < 3; ++c) px[k][c] = (c < C && inside) ? s[c * p.ch_stride] : (uint8_t)0;
            }
#pragma unroll
            for (int k = 0; k < GU; ++k) {
                const int q = q0 + k * gn;
                if (q < plane)
                    for (int c = 0; c < C; ++c) cur.st(c * plane + q, px[k][c]);
            }
        }
    }
    aug_sync<CL>();

    // ---- the sample's operations ------------------------------------------------------------
    const int32_t* rec = s_ops;
    for (int o = 0; o < p.n_ops; ++o, rec += 8) {
        const int kind = rec[0];
        const int r1 = rec[1];
        const float alpha = __int_as_float(r1);
        switch (kind) {
        case TB_AUG_AFFINE: {
            const uint32_t a0 = rec[1], a1 = rec[2], a2 = rec[3], a3 = rec[4], a4 = rec[5], a5 = rec[6];
            for (int q = gt; q < plane; q += gn) {
                const int y = (int)(((uint64_t)(uint32_t)q * wmagic) >> 40), x = q - y * W;      // q / W, exact for q < 2^20
                const int xin = (int)(a2 + (uint32_t)x * a0 + (uint32_t)y * a1) >> 16;
                const int yin = (int)(a5 + (uint32_t)x * a3 + (uint32_t)y * a4) >> 16;
                const bool ok = (unsigned)xin < (unsigned)W && (unsigned)yin < (unsigned)H;
                const int s = ok ? yin * W + xin : 0;
                for (int c = 0; c < C; ++c) alt.st(c * plane + q, ok ? (uint8_t)cur.ld(c * plane + s) : (uint8_t)0);
            }
            const AugBuf<CL> t = cur; cur = alt; alt = t;
            break;
        }
        case TB_AUG_BRIGHTNESS:
            aug_map<CL>(cur, total, gt, gn, [=](int, int v) { return aug_blend(0, v, alpha); });
            break;
        case TB_AUG_COLOR:
            if (C == 3)
                for (int q = gt; q < plane; q += gn) {
                    const int r = cur.ld(q), g = cur.ld(plane + q), bl = cur.ld(2 * plane + q);
                    const int l = aug_grey(r, g, bl);
                    cur.st(q, aug_blend(l, r, alpha));
                    cur.st(plane + q, aug_blend(l, g, alpha));
                    cur.st(2 * plane + q, aug_blend(l, bl, alpha));
                }
            break;
        case TB_AUG_CONTRAST: {
            if (tid == 0) s_sum = 0;
            __syncthreads();
            int part = 0;
            for (int q = gt; q < plane; q += gn)
                part += C == 3 ? aug_grey(cur.ld(q), cur.ld(plane + q), cur.ld(2 * plane + q)) : cur.ld(q);
            part = __reduce_add_sync(0xffffffffu, part);
            if ((tid & 31) == 0) atomicAdd(&s_sum, part);
            aug_sync<CL>();
            int sum = 0;
            for (int r = 0; r < S; ++r) sum += aug_peer<CL>(&s_sum, r);
            // int(sum / count + 0.5), Python float64
            const int mean = (int)__dadd_rn(__ddiv_rn((double)sum, (double)plane), 0.5);
            aug_map<CL>(cur, total, gt, gn, [=](int, int v) { return aug_blend(mean, v, alpha); });
            break;
        }
        case TB_AUG_SHARPNESS: {
            const float k1 = __fdiv_rn(1.0f, 13.0f), k5 = __fdiv_rn(5.0f, 13.0f);
            for (int q = gt; q < plane; q += gn) {
                const int y = (int)(((uint64_t)(uint32_t)q * wmagic) >> 40), x = q - y * W;      // q / W, exact for q < 2^20
                const bool interior = x >= 1 && x < W - 1 && y >= 1 && y < H - 1;
                for (int c = 0; c < C; ++c) {
                    const int at = c * plane + q;
                    const int v = cur.ld(at);
                    int deg = v;
                    if (interior) {
                        float ss = 0.5f;
#pragma unroll
                        for (int dy = 1; dy >= -1; --dy) {
                            const int row = at + dy * W;
                            const float mid = dy == 0 ? k5 : k1;
                            const float part = __fadd_rn(__fadd_rn(__fmul_rn((float)cur.ld(row - 1), k1),
                                                                   __fmul_rn((float)(dy == 0 ? v : cur.ld(row)), mid)),
                                                         __fmul_rn((float)cur.ld(row + 1), k1));
                            ss = __fadd_rn(ss, part);
                        }
                        deg = ss <= 0.0f ? 0 : (ss >= 255.0f ? 255 : (int)ss);
                    }
                    alt.st(at, aug_blend(deg, v, alpha));
                }
            }
            const AugBuf<CL> t = cur; cur = alt; alt = t;
            break;
        }
        case TB_AUG_POSTERIZE:
            aug_map<CL>(cur, total, gt, gn, [=](int, int v) { return (uint8_t)(v & r1); });
            break;
        case TB_AUG_SOLARIZE:
            aug_map<CL>(cur, total, gt, gn, [=](int, int v) { return (uint8_t)(v < r1 ? v : 255 - v); });
            break;
        case TB_AUG_INVERT:
            aug_map<CL>(cur, total, gt, gn, [=](int, int v) { return (uint8_t)(255 - v); });
            break;
        case TB_AUG_AUTOCONTRAST:
        case TB_AUG_EQUALIZE: {
            for (int i = tid; i < C * 256; i += nt) s_hist[i] = 0;
            if (tid < 3) { s_lo[tid] = 256; s_hi[tid] = -1; }
            __syncthreads();
            for (int i = gt * 4; i < total; i += gn * 4) {
                const uint32_t v = *cur.word(i);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (i + k < total) atomicAdd(&s_hist[aug_channel(i + k, plane) * 256 + ((v >> (8 * k)) & 0xffu)], 1);
            }
            aug_sync<CL>();
            const bool binner = tid < 256;          // threads 0..255 own one histogram bin each
            if (binner)
                for (int c = 0; c < C; ++c) {       // every CTA adds the shares up for itself
                    int n = 0;
                    for (int r = 0; r < S; ++r) n += aug_peer<CL>(&s_hist[c * 256 + tid], r);
                    s_full[c * 256 + tid] = n;
                    if (n) { atomicMin(&s_lo[c], tid); atomicMax(&s_hi[c], tid); }
                }
            __syncthreads();
            for (int c = 0; c < C; ++c) {
                const int lo = s_lo[c], hi = s_hi[c];
                int entry = tid;
                if (kind == TB_AUG_AUTOCONTRAST) {
                    if (binner && hi > lo) {      // ImageOps.autocontrast: scale = 255.0 / (hi - lo); offset = -lo * scale
                        const double scale = __ddiv_rn(255.0, (double)(hi - lo));
                        const double offset = __dmul_rn((double)(-lo), scale);
                        const double v = __dadd_rn(__dmul_rn((double)tid, scale), offset);
                        const int iv = (int)v;
                        entry = iv < 0 ? 0 : (iv > 255 ? 255 : iv);
                    }
                } else {
                    // ImageOps.equalize: step = (pixels - last non-empty bin) // 255;
                    // table[i] = (step // 2 + sum of bins below i) // step
                    const int h = binner ? s_full[c * 256 + tid] : 0;
                    int incl = h;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const int n = __shfl_up_sync(0xffffffffu, incl, d);
                        if ((tid & 31) >= d) incl += n;
                    }
                    __syncthreads();        // s_warp free again (previous channel read it)
                    if (binner && (tid & 31) == 31) s_warp[tid >> 5] = incl;
                    __syncthreads();
                    if (binner) {
                        int below = incl - h;
                        for (int w = 0; w < (tid >> 5); ++w) below += s_warp[w];
                        const int step = hi > lo ? (plane - s_full[c * 256 + hi]) / 255 : 0;
                        if (step > 0) {
                            const int n = (step / 2 + below) / step;
                            entry = n > 255 ? 255 : n;
                        }
                    }
                }
                if (binner) s_tab[c * 256 + tid] = (uint8_t)entry;
            }
            __syncthreads();
            aug_map<CL>(cur, total, gt, gn, [&](int i, int v) { return s_tab[aug_channel(i, plane) * 256 + v]; });
            break;
        }
        default:        // TB_AUG_IDENTITY and unknown codes
            break;
        }
        aug_sync<CL>();     // also keeps s_sum / s_hist alive until every CTA of the cluster has read them
    }

    // ---- table lookup, cast, planar store ---------------------------------------------------
    tb_launch_dependents();
    T* __restrict__ d = (T*)p.dst + (int64_t)b * total;
    if (sizeof(T) == 4 ? ((total & 3) == 0 && (((uintptr_t)p.dst) & 15) == 0)
                       : ((total & 7) == 0 && (((uintptr_t)p.dst) & 15) == 0)) {
        constexpr int PX = 16 / sizeof(T);
        for (int i = gt * PX; i < total; i += gn * PX) {
            __align__(16) T v[PX];
#pragma unroll
            for (int k = 0; k < PX; k += 4) {
                const uint32_t w = *cur.word(i + k);
#pragma unroll
                for (int j = 0; j < 4; ++j) v[k + j] = s_lut[aug_channel(i + k + j, plane) * 256 + ((w >> (8 * j)) & 0xffu)];
            }
            tb_st_cs_v4(d + i, *(const uint4*)v);
        }
    } else {
        for (int i = gt; i < total; i += gn) d[i] = s_lut[aug_channel(i, plane) * 256 + cur.ld(i)];
    }
    if constexpr (CL) cg::this_cluster().sync();        // no CTA leaves while its memory may still be read
}

template <typename T>
int launch_augment(const AugParams& p, int cluster, int threads, size_t smem, cudaStream_t s) {
    if (cluster == 1) {
        // static + dynamic shared memory beyond 48 KiB needs the opt-in (the kernel holds ~7 KiB statically)
        if (smem > 40 * 1024)
            TB_CUDA(cudaFuncSetAttribute(tb_augment_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        tb_augment_kernel<T, false><<<p.B, threads, smem, s>>>(p);
    } else {
        TB_CUDA(cudaFuncSetAttribute(tb_augment_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)p.B * cluster);
        cfg.blockDim = dim3(threads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)cluster;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        TB_CUDA(cudaLaunchKernelEx(&cfg, tb_augment_kernel<T, true>, p));
    }
    TB_LAUNCH_CHECK("tb_augment_kernel");
    return TB_OK;
}

// Whole-image nearest affine, global -> global: one thread per output pixel.
__global__ void __launch_bounds__(256) tb_affine_nearest_kernel(const uint8_t* __restrict__ src, int H, int W, int C, int64_t bs,
                                                                int64_t pitch, const int32_t* __restrict__ records,
                                                                uint8_t* __restrict__ dst, int blocks_per_image) {
    tb_grid_dependency_wait();
    const int b = blockIdx.x / blocks_per_image;
    const int q = (blockIdx.x - b * blocks_per_image) * 256 + threadIdx.x;
    if (q >= H * W) return;
    const int32_t* rec = records + (int64_t)b * 8;
    const int y = q / W, x = q - y * W;
    int xin = x, yin = y;
    if (rec[0] == TB_AUG_AFFINE) {
        const uint32_t a0 = rec[1], a1 = rec[2], a2 = rec[3], a3 = rec[4], a4 = rec[5], a5 = rec[6];
        xin = (int)(a2 + (uint32_t)x * a0 + (uint32_t)y * a1) >> 16;
        yin = (int)(a5 + (uint32_t)x * a3 + (uint32_t)y * a4) >> 16;
    }
    const bool ok = (unsigned)xin < (unsigned)W && (unsigned)yin < (unsigned)H;
    const uint8_t* s = src + (int64_t)b * bs + (int64_t)(ok ? yin : 0) * pitch + (int64_t)(ok ? xin : 0) * C;
    uint8_t* d = dst + ((int64_t)b * H * W + q) * C;
    for (int c = 0; c < C; ++c) d[c] = ok ? s[c] : (uint8_t)0;
}

}  // namespace

extern "C" int tb_affine_nearest_u8_nhwc(const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride,
                                         int64_t src_row_pitch, const int32_t* records, uint8_t* dst, tb_stream_t stream) {
    TB_REQUIRE(B >= 0 && H >= 1 && W >= 1 && H < 32768 && W < 32768, "tb_affine_nearest: bad shape B=%d H=%d W=%d", B, H, W);
    TB_REQUIRE(C >= 1 && C <= 4, "tb_affine_nearest: C=%d not in [1,4]", C);
    TB_REQUIRE(src_row_pitch >= (int64_t)W * C && src_batch_stride >= 0, "tb_affine_nearest: bad strides");
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && dst && records, "tb_affine_nearest: NULL argument");
    const int64_t bpi = ((int64_t)H * W + 255) / 256;
    TB_REQUIRE(bpi * B < ((int64_t)1 << 31), "tb_affine_nearest: grid too large");
    tb_affine_nearest_kernel<<<(unsigned)(bpi * B), 256, 0, (cudaStream_t)stream>>>(src, H, W, C, src_batch_stride, src_row_pitch,
                                                                                  records, dst, (int)bpi);
    TB_LAUNCH_CHECK("tb_affine_nearest_kernel");
    return TB_OK;
}

extern "C" int tb_augment_u8_nhwc_to_nchw(const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride,
                                          int64_t src_row_pitch, void* dst, int dst_dtype, int out_h, int out_w,
                                          const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
                                          int pad_y, int pad_x, int pad_mode, const int32_t* ops, int n_ops,
                                          const void* lut, int flags, tb_stream_t stream) {
    TB_REQUIRE(B >= 0 && H >= 1 && W >= 1, "tb_augment: bad source shape B=%d H=%d W=%d", B, H, W);
    TB_REQUIRE(C == 1 || C == 3, "tb_augment: C=%d not in {1,3} (PIL modes L and RGB)", C);
    TB_REQUIRE(out_h >= 1 && out_w >= 1, "tb_augment: bad output size %dx%d", out_h, out_w);
    TB_REQUIRE(out_h < 32768 && out_w < 32768, "tb_augment: output beyond the 16.16 fixed-point range");
    TB_REQUIRE(tb_dtype_size(dst_dtype) != 0, "tb_augment: unknown dst_dtype %d", dst_dtype);
    TB_REQUIRE(pad_mode >= TB_PAD_NONE && pad_mode <= TB_PAD_SYMMETRIC, "tb_augment: unknown pad_mode %d", pad_mode);
    TB_REQUIRE(pad_y >= 0 && pad_x >= 0, "tb_augment: negative padding");
    if (pad_mode == TB_PAD_REFLECT)
        TB_REQUIRE(pad_y <= H - 1 && pad_x <= W - 1, "tb_augment: reflect padding must be smaller than the image");
    if (pad_mode == TB_PAD_SYMMETRIC)
        TB_REQUIRE(pad_y <= H && pad_x <= W, "tb_augment: symmetric padding must not exceed the image");
    const bool planar = (flags & TB_FLAG_SRC_PLANAR) != 0;
    TB_REQUIRE(src_row_pitch >= (int64_t)W * (planar ? 1 : C), "tb_augment: src_row_pitch %lld shorter than a row", (long long)src_row_pitch);
    TB_REQUIRE(src_batch_stride >= 0, "tb_augment: negative src_batch_stride");
    TB_REQUIRE(n_ops >= 0 && n_ops <= TB_AUG_MAX_OPS, "tb_augment: n_ops=%d not in [0,%d]", n_ops, TB_AUG_MAX_OPS);
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && dst && lut, "tb_augment: NULL src/dst/lut");
    TB_REQUIRE(ops || n_ops == 0, "tb_augment: NULL ops");
    TB_REQUIRE(((uintptr_t)dst % tb_dtype_size(dst_dtype)) == 0, "tb_augment: dst misaligned for its dtype");

    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;

    AugParams p;
    p.src = src; p.B = B; p.H = H; p.W = W; p.C = C; p.bs = src_batch_stride; p.pitch = src_row_pitch;
    p.dst = dst; p.out_h = out_h; p.out_w = out_w; p.top = crop_top; p.left = crop_left; p.flip = flip;
    p.pad_y = pad_y; p.pad_x = pad_x; p.pad_mode = pad_mode; p.ops = ops; p.n_ops = n_ops; p.lut = lut;
    p.px_stride = planar ? 1 : C;
    p.ch_stride = planar ? (int64_t)H * src_row_pitch : 1;
    const int64_t plane = (int64_t)out_h * out_w;
    // The image lives in shared memory, twice (operations that move pixels write the other copy):
    // in one CTA when it fits, else spread over a cluster of 2, 4 or 8 CTAs (distributed shared memory).
    const int64_t reserve = 10 * 1024;      // static shared memory of the kernel + the system's share
    int cluster = (flags >> 8) & 0xf;       // development knob: force a cluster size
    int64_t buf = 0, smem = 0;
    // Small chunks win: shared memory per CTA decides how many warps an SM holds (224 x 224 RGB:
    // 2 CTAs of 150 KB = 8 warps per SM, 8 CTAs of 38 KB = 40), so an image takes the smallest
    // cluster whose chunks are at most kAugChunk bytes, or 8 CTAs.
    const bool forced = cluster != 0;
    for (int c = forced ? cluster : 1; c <= 8; c *= 2) {
        buf = (((plane * C + c - 1) / c) + 15) & ~(int64_t)15;
        smem = 2 * buf + (int64_t)C * 256 * tb_dtype_size(dst_dtype);
        const bool fits = smem + reserve <= dev.max_smem_optin;
        if (forced) { if (!fits) cluster = 0; break; }
        if (fits && (c == 8 || buf <= kAugChunk)) { cluster = c; break; }
    }
    if (cluster == 0 || smem + reserve > dev.max_smem_optin || plane * C >= ((int64_t)1 << 20))
        return tb_fail(TB_ENOTSUP, "tb_augment: a %dx%dx%d image does not fit twice in the shared memory of a cluster of 8 SMs",
                       out_h, out_w, C);
    TB_REQUIRE(cluster == 1 || cluster == 2 || cluster == 4 || cluster == 8, "tb_augment: cluster size %d", cluster);
    TB_REQUIRE((int64_t)B * cluster < ((int64_t)1 << 31), "tb_augment: grid too large");
    p.plane = (int)plane;
    p.buf_bytes = (int)buf;
    p.magic = ((((uint64_t)1) << 40) + (uint64_t)buf - 1) / (uint64_t)buf;
    p.wmagic = ((((uint64_t)1) << 40) + (uint64_t)out_w - 1) / (uint64_t)out_w;
    cudaStream_t s = (cudaStream_t)stream;
    // few images: more threads per image shorten every pass (a CTA walks its image pass by pass)
    int threads = kAugThreads;
    const int64_t ctas = (int64_t)B * cluster;
    if (ctas <= 2 * (int64_t)dev.sm_count) threads = 512;       // measured (ncu, 256 CIFAR images): 256 threads 28.8 us, 512 22.8, 1024 28.4
    while (threads > kAugThreads && (int64_t)threads * 4 > ((plane * C + cluster - 1) / cluster) * 2) threads >>= 1;     // >= 2 bytes per thread and pass
    if ((flags >> 12) & 0xf) threads = 64 << ((flags >> 12) & 0xf);       // development knob: 128 .. 1024
    if (threads < kAugThreads) threads = kAugThreads;
    if (threads > kAugMaxThreads) threads = kAugMaxThreads;
    return dst_dtype == TB_F32 ? launch_augment<float>(p, cluster, threads, (size_t)smem, s)
                               : launch_augment<uint16_t>(p, cluster, threads, (size_t)smem, s);
}
