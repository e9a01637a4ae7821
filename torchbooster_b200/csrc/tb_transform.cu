// Fused pad -> crop -> horizontal flip -> table lookup -> cast, uint8 NHWC -> NCHW (sm_100a).
//
// Replaces the per-sample torchvision chain of the reference's examples
// (examples/img_cls/resnet/resnet.py:94-103: RandomCrop(32, 4, "reflect"), RandomHorizontalFlip,
// ToTensor, Normalize; collate at torchbooster/config.py:371-379) with one launch per batch.
//
// Two kernels, same index map (DESIGN.md "transform"):
//   * tb_xf_generic_kernel : one gather per output pixel, any C in [1,4], any geometry / padding.
//   * tb_xf_stream_kernel  : the HBM-bound path for C == 3.  One persistent CTA per SM.  Every
//     warp streams its own tiles (a few crop rows of one image): it issues 1-D bulk async copies
//     (TMA engine, cp.async.bulk -> SASS UBLKCP) of the rows into its private two-stage
//     shared-memory ring, waits on its own mbarrier, de-interleaves / flips out of shared
//     memory, looks every byte up in a lane-replicated table and writes 16-byte streaming stores
//     to the three planes.  Warps never wait for each other, so they drift apart and cover each
//     other's latencies.  Tiles whose crop window leaves the image (padding) take the gather
//     route inside the same launch, so any parameter values are safe.
// The arithmetic is a table lookup on purpose: see tb_lut.cu.
#include "tb_common.cuh"

namespace {

struct XfParams {
    const uint8_t* src;
    const int64_t* offsets;     // per-image byte offsets from src (records mode), or NULL: image b at b * bs
    int B, H, W, C;
    int64_t bs, pitch;
    void* dst;
    int out_h, out_w;
    const int32_t* top;
    const int32_t* left;
    const uint8_t* flip;
    int pad_y, pad_x, pad_mode;
    const void* lut;
};

__device__ __forceinline__ const uint8_t* tb_image(const XfParams& p, int b) {
    return p.src + (p.offsets ? p.offsets[b] : (int64_t)b * p.bs);
}

// Source coordinate of padded coordinate i for an axis of length n; `inside` cleared when the
// element is fill (0).  numpy.pad semantics (torchvision/transforms/_functional_pil.py:200-223).
__device__ __forceinline__ int tb_fold(int i, int n, int mode, bool& inside) {
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

// ------------------------------------------------------------------------------------------
// generic: one thread per output pixel, all channels
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) tb_xf_generic_kernel(XfParams p, int blocks_per_image) {
    __shared__ T s_lut[4 * 256];
    tb_launch_dependents();
    tb_grid_dependency_wait();
    const T* __restrict__ glut = (const T*)p.lut;
    for (int i = threadIdx.x; i < p.C * 256; i += blockDim.x) s_lut[i] = glut[i];
    __syncthreads();

    const int b = blockIdx.x / blocks_per_image;
    const int pix = (blockIdx.x - b * blocks_per_image) * blockDim.x + threadIdx.x;
    if (pix >= p.out_h * p.out_w) return;
    const int y = pix / p.out_w;
    const int x = pix - y * p.out_w;
    const int top = (p.top ? p.top[b] : 0) - p.pad_y;
    const int left = (p.left ? p.left[b] : 0) - p.pad_x;
    const bool flip = p.flip ? (p.flip[b] != 0) : false;
    const int xs = flip ? p.out_w - 1 - x : x;
    bool inside = true;
    const int u = tb_fold(top + y, p.H, p.pad_mode, inside);
    const int v = tb_fold(left + xs, p.W, p.pad_mode, inside);
    const uint8_t* __restrict__ s = tb_image(p, b) + (int64_t)u * p.pitch + (int64_t)v * p.C;
    T* __restrict__ d = (T*)p.dst + ((int64_t)b * p.C * p.out_h + y) * p.out_w + x;
    const int64_t plane = (int64_t)p.out_h * p.out_w;
    for (int c = 0; c < p.C; ++c) {
        const int e = inside ? (int)s[c] : 0;
        d[c * plane] = s_lut[c * 256 + e];
    }
}

// ------------------------------------------------------------------------------------------
// streaming: C == 3, persistent, per-warp bulk-copy rings + lane-replicated table
// ------------------------------------------------------------------------------------------
// Shared-window map (byte addresses as seen by ld.shared).  A table row is 256 bytes (64 words)
// per byte value, so the byte to look up drops into address bits 8-15 with ONE PRMT, and lane l
// always reads word l (or 32 + l) of its row, i.e. its own bank: no bank conflicts whatever the
// 32 byte values are.
//   16-bit outputs (bf16 / fp16):
//     [s0, 0x20000)       plain table copy (3 KiB), per-warp rings, mbarriers
//     [0x20000, 0x30000)  rows: words 0-31 = {c0 | c1 << 16} per lane, words 32-63 = {c2 | c2 << 16}
//   32-bit outputs (fp32):
//     [s0, 0x10000)       plain table copy, per-warp rings, mbarriers
//     [0x10000, 0x20000)  rows: words 0-31 = c0, words 32-63 = c1
//     [0x20000, 0x30000)  rows: words 0-31 = c2
constexpr uint32_t kXfSmemEnd = 0x30000u;
constexpr int kXfHostCtas = 8;       // CTAs of a launch whose source is pinned host memory

__device__ __forceinline__ uint32_t tb_lds32(uint32_t addr) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ void tb_lds64(uint32_t addr, uint32_t& lo, uint32_t& hi) {
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(addr));
}

template <typename T>
struct Xf;   // per output width: pixels per 16-byte store, table layout, packing
template <>
struct Xf<float> {
    static constexpr int PX = 4;
    static constexpr uint32_t kLow = 0x10000u;    // end of the ring region
    static constexpr uint32_t kTab = 0x10000u;    // first table
    __device__ static __forceinline__ uint32_t widen(float f) { return __float_as_uint(f); }
    // 16 x 16-byte stores per row: words 0-31 one channel, words 32-63 the next
    __device__ static __forceinline__ void fill(uint32_t* tab, const uint32_t* plain, int tid, int nthreads) {
        uint4* t4 = reinterpret_cast<uint4*>(tab);
        for (int i = tid; i < 2 * 256 * 16; i += nthreads) {
            const int t = i >> 12, e = (i >> 4) & 255, c = t * 2 + ((i >> 3) & 1);
            if (c < 3) {
                const uint32_t v = plain[c * 256 + e];
                t4[i] = make_uint4(v, v, v, v);
            }
        }
    }
    __device__ static __forceinline__ uint32_t base(int c, int lane) {
        return kTab + (c == 2 ? 0x10000u : 0u) + (c == 1 ? 128u : 0u) + lane * 4;
    }
    __device__ static __forceinline__ uint4 pack(const uint32_t* v, int) { return make_uint4(v[0], v[1], v[2], v[3]); }
    __device__ static __forceinline__ float one(uint32_t w, int) { return __uint_as_float(w); }
};
template <>
struct Xf<uint16_t> {   // bf16 and fp16: the table already holds the rounded 16-bit pattern
    static constexpr int PX = 8;
    static constexpr uint32_t kLow = 0x20000u;
    static constexpr uint32_t kTab = 0x20000u;
    __device__ static __forceinline__ uint32_t widen(uint16_t h) { return h; }
    __device__ static __forceinline__ void fill(uint32_t* tab, const uint32_t* plain, int tid, int nthreads) {
        uint4* t4 = reinterpret_cast<uint4*>(tab);
        for (int i = tid; i < 256 * 16; i += nthreads) {
            const int e = i >> 4;
            const uint32_t v = (i & 8) ? plain[512 + e] * 0x10001u : (plain[e] | (plain[256 + e] << 16));
            t4[i] = make_uint4(v, v, v, v);
        }
    }
    __device__ static __forceinline__ uint32_t base(int c, int lane) { return kTab + (c == 2 ? 128u : 0u) + lane * 4; }
    __device__ static __forceinline__ uint4 pack(const uint32_t* v, int c) {
        const uint32_t sel = c == 1 ? 0x7632u : 0x5410u;   // channel 1 lives in the high halves
        return make_uint4(__byte_perm(v[0], v[1], sel), __byte_perm(v[2], v[3], sel), __byte_perm(v[4], v[5], sel),
                          __byte_perm(v[6], v[7], sel));
    }
    __device__ static __forceinline__ uint16_t one(uint32_t w, int c) { return (uint16_t)(c == 1 ? w >> 16 : w); }
};

struct XfGeom {
    int rows;                 // output rows per warp tile (<= 32: one bulk copy per lane)
    int rowbuf;               // bytes per staged row (multiple of 16)
    int contiguous;           // crop rows are whole source rows back to back: one bulk copy per tile
    int tiles_per_image, total_tiles;
    int gpr;                  // 16-byte output groups ("units") per row
    int step_r, step_g;       // 32 units further = (step_r rows, step_g groups) further
    int tile_db, tile_dt;     // stride of the tile walk (all warps of the grid) as (images, tiles)
    int stage_bytes;
    uint32_t dyn_bytes;
};

template <typename T>
__global__ void __launch_bounds__(768, 1) tb_xf_stream_kernel(XfParams p, XfGeom g) {
    using X = Xf<T>;
    constexpr int PX = X::PX;
    constexpr int NB = 3 * PX;            // source bytes per unit
    constexpr int NW = NB / 4;            // aligned words once shifted
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t s0 = tb_smem_addr(smem);
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int nwarps = (int)blockDim.x >> 5;
    const uint32_t plain_addr = (s0 + 15u) & ~15u;
    const uint32_t ring0 = (plain_addr + 3072u + 127u) & ~127u;
    const uint32_t bars_addr = ring0 + (uint32_t)(nwarps * 2 * g.stage_bytes);
    // the launch must have provided the window this map assumes
    if (bars_addr + (uint32_t)nwarps * 16u > X::kLow || s0 + g.dyn_bytes < kXfSmemEnd) asm volatile("trap;");
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (bars_addr - s0)) + warp * 2;
    uint32_t* plain = reinterpret_cast<uint32_t*>(smem + (plain_addr - s0));
    uint32_t* tab = reinterpret_cast<uint32_t*>(smem + (X::kTab - s0));
    const uint32_t my_ring = ring0 + (uint32_t)(warp * 2 * g.stage_bytes);
    uint8_t* my_ring_ptr = smem + (my_ring - s0);

    if (lane == 0) {
        tb_mbar_init(&full[0], 1);
        tb_mbar_init(&full[1], 1);
        tb_mbar_fence_init();
    }
    __syncwarp();
    // everything above touches no global memory: with programmatic dependent launch it overlaps
    // the tail of the previous kernel in the stream
    tb_launch_dependents();
    tb_grid_dependency_wait();

    // two cursors over this warp's tiles t = gw, gw + stride, ...: the load cursor runs two tiles
    // ahead of the compute cursor
    const int gw = blockIdx.x * nwarps + warp;
    int tp = gw, bp = gw / g.tiles_per_image, tip = gw - bp * g.tiles_per_image;
    int tc = tp, bc = bp, tic = tip;
    const int tstride = g.tile_db * g.tiles_per_image + g.tile_dt;
    auto advance = [&](int& t, int& b, int& ti) {
        t += tstride;
        b += g.tile_db;
        ti += g.tile_dt;
        if (ti >= g.tiles_per_image) { ti -= g.tiles_per_image; ++b; }
    };
    // Crop parameters of the tile a stage holds, read ONCE (when its copies are issued) and kept for the compute side:
    // a tile costs one round trip for (top, left, flip) instead of two, and none on the path that waits for its data.
    int top_s0 = 0, left_s0 = 0, top_s1 = 0, left_s1 = 0;
    bool flip_s0 = false, flip_s1 = false;
    auto load_params = [&](int b, int& top, int& left, bool& flip) {
        top = (p.top ? p.top[b] : 0) - p.pad_y;
        left = (p.left ? p.left[b] : 0) - p.pad_x;
        flip = p.flip ? (p.flip[b] != 0) : false;
    };
    auto issue_tile = [&](int stage, int t, int b, int ti, int top, int left) {
        // bulk copies of tile (b, ti) into `stage`; nothing is armed for tiles that leave the image
        if (t < g.total_tiles) {
            const int y0 = ti * g.rows;
            const int nrows = min(g.rows, p.out_h - y0);
            const bool inside = top >= 0 && top + p.out_h <= p.H && left >= 0 && left + p.out_w <= p.W;
            if (inside) {
                const int boff = left * 3;
                const uint8_t* gsrc = tb_image(p, b) + (int64_t)(top + y0) * p.pitch + (boff & ~15);
                uint8_t* sdst = my_ring_ptr + stage * g.stage_bytes;
                if (g.contiguous) {
                    if (lane == 0) {
                        const uint32_t len = (uint32_t)nrows * (uint32_t)g.rowbuf;
                        tb_mbar_arrive_expect_tx(&full[stage], len);
                        tb_bulk_g2s(sdst, gsrc, len, &full[stage]);
                    }
                } else {
                    // 16-byte aligned superset of the crop row
                    const uint32_t len = (uint32_t)(((boff & 15) + p.out_w * 3 + 15) & ~15);
                    if (lane == 0) tb_mbar_arrive_expect_tx(&full[stage], len * (uint32_t)nrows);
                    __syncwarp();
                    if (lane < nrows)
                        tb_bulk_g2s(sdst + (size_t)lane * g.rowbuf, gsrc + (int64_t)lane * p.pitch, len, &full[stage]);
                }
            }
        }
    };
    auto issue = [&](int stage) {
        int top = 0, left = 0;
        bool flip = false;
        if (tp < g.total_tiles) load_params(bp, top, left, flip);
        issue_tile(stage, tp, bp, tip, top, left);
        if (stage) { top_s1 = top; left_s1 = left; flip_s1 = flip; }
        else { top_s0 = top; left_s0 = left; flip_s0 = flip; }
        advance(tp, bp, tip);
    };
    // Start-up: everything the CTA needs from global memory is asked for at once -- the table values (at most four per
    // thread, into registers) and the parameters of the first TWO tiles -- so the launch pays one round trip, not the
    // four in a row it used to (top, left of tile 0; of tile 1; then the table: 5 us before the first tile was touched).
    T lutv[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int i = tid + q * (int)blockDim.x;
        lutv[q] = i < 768 ? ((const T*)p.lut)[i] : T(0);
    }
    {
        const int tA = tp, bA = bp, tiA = tip;
        advance(tp, bp, tip);
        const int tB = tp, bB = bp, tiB = tip;
        advance(tp, bp, tip);
        if (tA < g.total_tiles) load_params(bA, top_s0, left_s0, flip_s0);
        if (tB < g.total_tiles) load_params(bB, top_s1, left_s1, flip_s1);
        issue_tile(0, tA, bA, tiA, top_s0, left_s0);
        issue_tile(1, tB, bB, tiB, top_s1, left_s1);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int i = tid + q * (int)blockDim.x;
        if (i < 768) plain[i] = X::widen(lutv[q]);
    }
    for (int i = tid + 4 * (int)blockDim.x; i < 768; i += blockDim.x) plain[i] = X::widen(((const T*)p.lut)[i]);   // (fewer than 192 threads)
    __syncthreads();
    X::fill(tab, plain, tid, blockDim.x);
    __syncthreads();

    const uint32_t base0 = X::base(0, lane), base1 = X::base(1, lane), base2 = X::base(2, lane);
    const int64_t plane = (int64_t)p.out_h * p.out_w;
    const int ur0 = lane / g.gpr;                                   // (row, group) of this lane's first unit
    const int ug0 = lane - ur0 * g.gpr;

    uint32_t phases = 0;     // bit s = parity to wait for on stage s
    for (int it = 0; tc < g.total_tiles; ++it) {
        const int stage = it & 1;
        const int y0 = tic * g.rows;
        const int nrows = min(g.rows, p.out_h - y0);
        const int top = stage ? top_s1 : top_s0;
        const int left = stage ? left_s1 : left_s0;
        const bool flip = stage ? flip_s1 : flip_s0;
        const bool inside = top >= 0 && top + p.out_h <= p.H && left >= 0 && left + p.out_w <= p.W;
        T* __restrict__ dimg = (T*)p.dst + (int64_t)bc * 3 * plane;

        if (inside) {
            tb_mbar_wait(&full[stage], (phases >> stage) & 1u);
            phases ^= 1u << stage;
            const uint32_t sbase = my_ring + (uint32_t)(stage * g.stage_bytes);
            const int units = nrows * g.gpr;
            // first source byte of group 0 and the signed byte step from one group to the next
            const int start0 = ((left * 3) & 15) + (flip ? 3 * (p.out_w - PX) : 0);
            const int gstep = flip ? -3 * PX : 3 * PX;
            int r = ur0, gq = ug0;
            T* d = dimg + (int64_t)y0 * p.out_w + (int64_t)lane * PX;   // a tile is contiguous per plane
            for (int u = lane; u < units; u += 32, d += 32 * PX) {
                // NB source bytes from `start`: 8-byte loads of the aligned window (lanes 24 bytes
                // apart never share a bank within a half-warp), then a word select + funnel shift.
                const int start = start0 + gq * gstep;
                const uint32_t wa = sbase + (uint32_t)(r * g.rowbuf + (start & ~7));
                const bool hi4 = (start & 4) != 0;
                const int sh = (start & 3) * 8;
                constexpr int NDW = (NW + 3) / 2;     // 8-byte words covering (start & 7) + NB bytes
                uint32_t w[2 * NDW];
#pragma unroll
                for (int i = 0; i < NDW; ++i) tb_lds64(wa + 8 * i, w[2 * i], w[2 * i + 1]);
                uint32_t q[NW];
#pragma unroll
                for (int i = 0; i < NW; ++i)
                    q[i] = __funnelshift_r(hi4 ? w[i + 1] : w[i], hi4 ? w[i + 2] : w[i + 1], sh);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t base = c == 0 ? base0 : (c == 1 ? base1 : base2);
                    uint32_t v[PX];
                    if (flip) {     // the flip only changes which source pixel feeds output slot j
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const int bi = 3 * (PX - 1 - j) + c;
                            v[j] = tb_lds32(__byte_perm(base, q[bi >> 2], 0x3240 | ((bi & 3) << 4)));
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const int bi = 3 * j + c;
                            v[j] = tb_lds32(__byte_perm(base, q[bi >> 2], 0x3240 | ((bi & 3) << 4)));
                        }
                    }
                    tb_st_cs_v4(d + c * plane, X::pack(v, c));
                }
                r += g.step_r;
                gq += g.step_g;
                if (gq >= g.gpr) { gq -= g.gpr; ++r; }
            }
        } else {
            // crop window touches padding: gather route for this tile
            const int npix = nrows * p.out_w;
            for (int i = lane; i < npix; i += 32) {
                const int r = i / p.out_w;
                const int x = i - r * p.out_w;
                const int xs = flip ? p.out_w - 1 - x : x;
                bool in = true;
                const int uu = tb_fold(top + y0 + r, p.H, p.pad_mode, in);
                const int vv = tb_fold(left + xs, p.W, p.pad_mode, in);
                const uint8_t* sp = tb_image(p, bc) + (int64_t)uu * p.pitch + (int64_t)vv * 3;
                T* dd = dimg + (int64_t)(y0 + r) * p.out_w + x;
                const uint32_t e0 = in ? sp[0] : 0u, e1 = in ? sp[1] : 0u, e2 = in ? sp[2] : 0u;
                dd[0] = X::one(tb_lds32(base0 + (e0 << 8)), 0);
                dd[plane] = X::one(tb_lds32(base1 + (e1 << 8)), 1);
                dd[2 * plane] = X::one(tb_lds32(base2 + (e2 << 8)), 2);
            }
        }
        // this stage is drained by the whole warp: refill it with the tile two steps ahead
        __syncwarp();
        tb_fence_proxy_async();
        issue(stage);
        advance(tc, bc, tic);
    }
}

template <typename T>
int launch_generic(const XfParams& p, cudaStream_t s, bool pdl) {
    // a block copies the table into shared memory before its pixels: large blocks (a whole 32 x 32 image each) keep
    // that fixed cost per pixel low; small outputs keep small blocks
    const int64_t pixels = (int64_t)p.out_h * p.out_w;
    const int threads = pixels >= 1024 ? 1024 : (pixels >= 512 ? 512 : 256);
    const int64_t bpi = (pixels + threads - 1) / threads;
    const int64_t blocks = bpi * p.B;
    TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_transform: grid too large");
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3((unsigned)threads);
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, tb_xf_generic_kernel<T>, p, (int)bpi));
    TB_LAUNCH_CHECK("tb_xf_generic_kernel");
    return TB_OK;
}

template <typename T>
int launch_stream(const XfParams& p, const TbDevice& dev, cudaStream_t s, int force_rows, int force_warps,
                  int max_ctas, bool pdl, bool* taken) {
    using X = Xf<T>;
    constexpr int PX = X::PX;
    *taken = false;
    if (dev.max_smem_optin < (int)kXfSmemEnd + 2048) return TB_OK;
    XfGeom g;
    // Whole-row crops of a dense source land as one copy per tile; otherwise one copy per row
    // covering the 16-byte aligned superset of the crop row.
    g.contiguous = (p.W == p.out_w && p.pitch == (int64_t)p.W * 3) ? 1 : 0;
    g.rowbuf = g.contiguous ? (int)p.pitch : ((p.out_w * 3 + 15 + 15) & ~15);
    g.gpr = p.out_w / PX;
    // dynamic shared memory starts within the first 2 KiB of the window (1 KiB is reserved by the
    // system and this kernel has no static shared memory); the kernel traps if that ever changes
    const int avail = (int)X::kLow - 2048 - 3072 - 256;
    // Rows per warp tile R and warps per CTA W.  Issue slots per image do not depend on W, so:
    // fewest slots per image (a fixed cost per tile plus one unit per 32-lane round), with a mild
    // preference for more resident warps.
    int best_rows = 0, best_warps = 0;
    double best = -1.0;
    for (int warps = 6; warps <= 24; ++warps) {
        if (force_warps && warps != force_warps) continue;
        const int budget = (avail - 16 * warps) / warps;
        int max_rows = budget / (2 * g.rowbuf);
        if (max_rows > 32) max_rows = 32;
        if (max_rows > p.out_h) max_rows = p.out_h;
        for (int rows = 1; rows <= max_rows; ++rows) {
            if (force_rows && rows != force_rows) continue;
            const int full = p.out_h / rows, tail = p.out_h - full * rows;
            auto tile_cost = [&](int r) { return ((r * g.gpr + 31) / 32) * 110.0 + 90.0; };
            double cost = full * tile_cost(rows) + (tail ? tile_cost(tail) : 0.0);
            cost *= 1.0 + 6.0 / warps;
            if (best < 0 || cost < best) { best = cost; best_rows = rows; best_warps = warps; }
        }
    }
    if (best < 0) return TB_OK;     // a single row does not fit a warp's ring: gather path
    g.rows = best_rows;
    const int warps = best_warps;
    g.stage_bytes = g.rows * g.rowbuf;
    g.tiles_per_image = (p.out_h + g.rows - 1) / g.rows;
    const int64_t total = (int64_t)g.tiles_per_image * p.B;
    if (total >= (int64_t)1 << 30) return TB_OK;
    g.total_tiles = (int)total;
    g.step_r = 32 / g.gpr;
    g.step_g = 32 % g.gpr;
    int64_t grid = (total + warps - 1) / warps;
    if (grid > dev.sm_count) grid = dev.sm_count;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    const int64_t stride = grid * warps;
    g.tile_db = (int)(stride / g.tiles_per_image);
    g.tile_dt = (int)(stride % g.tiles_per_image);
    g.dyn_bytes = kXfSmemEnd;
    static int attr_set[64] = {};            // per device, atomic: racing first calls set the same attribute
    if (!__atomic_load_n(&attr_set[dev.device], __ATOMIC_ACQUIRE)) {
        TB_CUDA(cudaFuncSetAttribute(tb_xf_stream_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     dev.max_smem_optin));
        __atomic_store_n(&attr_set[dev.device], 1, __ATOMIC_RELEASE);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)warps * 32);
    cfg.dynamicSmemBytes = g.dyn_bytes;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, tb_xf_stream_kernel<T>, p, g));
    TB_LAUNCH_CHECK("tb_xf_stream_kernel");
    *taken = true;
    return TB_OK;
}

// Chance that a uniformly drawn crop window lies inside the unpadded image along one axis.
double inside_fraction(int n, int pad, int out) {
    const int positions = n + 2 * pad - out + 1;      // RandomCrop draws from [0, positions)
    const int inside = n - out + 1;
    if (positions <= 0 || inside <= 0) return 0.0;
    return (double)inside / positions;
}

}  // namespace

static int tb_transform_impl(const uint8_t* src, const int64_t* offsets, int B, int H, int W, int C,
                             int64_t src_batch_stride, int64_t src_row_pitch, void* dst, int dst_dtype, int out_h,
                             int out_w, const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
                             int pad_y, int pad_x, int pad_mode, const void* lut, int flags, tb_stream_t stream) {
    TB_REQUIRE(B >= 0 && H >= 1 && W >= 1, "tb_transform: bad source shape B=%d H=%d W=%d", B, H, W);
    TB_REQUIRE(C >= 1 && C <= 4, "tb_transform: C=%d not in [1,4]", C);
    TB_REQUIRE(out_h >= 1 && out_w >= 1, "tb_transform: bad output size %dx%d", out_h, out_w);
    TB_REQUIRE(tb_dtype_size(dst_dtype) != 0, "tb_transform: unknown dst_dtype %d", dst_dtype);
    TB_REQUIRE(pad_mode >= TB_PAD_NONE && pad_mode <= TB_PAD_SYMMETRIC, "tb_transform: unknown pad_mode %d", pad_mode);
    TB_REQUIRE(pad_y >= 0 && pad_x >= 0, "tb_transform: negative padding");
    if (pad_mode == TB_PAD_REFLECT)
        TB_REQUIRE(pad_y <= H - 1 && pad_x <= W - 1, "tb_transform: reflect padding must be smaller than the image");
    if (pad_mode == TB_PAD_SYMMETRIC)
        TB_REQUIRE(pad_y <= H && pad_x <= W, "tb_transform: symmetric padding must not exceed the image");
    TB_REQUIRE(src_row_pitch >= (int64_t)W * C, "tb_transform: src_row_pitch %lld < W*C", (long long)src_row_pitch);
    TB_REQUIRE(src_batch_stride >= 0, "tb_transform: negative src_batch_stride");
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && dst && lut, "tb_transform: NULL src/dst/lut");
    TB_REQUIRE(((uintptr_t)dst % tb_dtype_size(dst_dtype)) == 0, "tb_transform: dst misaligned for its dtype");

    XfParams p;
    p.src = src; p.offsets = offsets; p.B = B; p.H = H; p.W = W; p.C = C; p.bs = src_batch_stride; p.pitch = src_row_pitch;
    p.dst = dst; p.out_h = out_h; p.out_w = out_w;
    p.top = crop_top; p.left = crop_left; p.flip = flip;
    p.pad_y = pad_y; p.pad_x = pad_x; p.pad_mode = pad_mode; p.lut = lut;
    cudaStream_t s = (cudaStream_t)stream;

    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;

    const int px = 16 / tb_dtype_size(dst_dtype);
    bool stream_ok = !(flags & TB_FLAG_FORCE_GENERIC) && C == 3 && (out_w % px) == 0 &&
                     ((uintptr_t)dst % 16) == 0 && ((uintptr_t)src % 16) == 0 && (src_row_pitch % 16) == 0 &&
                     (src_batch_stride % 16) == 0;
    // With padding most random windows leave the image (CIFAR: 32 from 32 + 2*4 -> 1 in 81 stays
    // inside) and would take the gather route anyway; the plain gather kernel is then the better
    // launch.  TB_FLAG_FORCE_STREAM overrides (tests).
    if (stream_ok && !(flags & TB_FLAG_FORCE_STREAM) && (crop_top || crop_left) &&
        inside_fraction(H, pad_y, out_h) * inside_fraction(W, pad_x, out_w) < 0.5)
        stream_ok = false;
    if (stream_ok) {
        bool taken = false;
        // development knobs: bits 8-15 CTA limit, bits 16-23 warps per CTA, bits 24-31 rows per warp tile
        const int fw = (flags >> 16) & 0xff, fr = (flags >> 24) & 0xff;
        const bool pdl = !(flags & TB_FLAG_NO_PDL);
        // A batch that lives in pinned HOST memory is read in place: the bulk copies go over PCIe
        // and fetch only the crop rows.  PCIe needs little parallelism (4 CTAs already saturate it;
        // whole-row requests were measured slower than row windows), so such a launch takes a few
        // SMs only and leaves the rest of the GPU to whatever else is running.
        int max_ctas = (flags >> 8) & 0xff;
        if (max_ctas == 0) {
            cudaPointerAttributes pa;
            if (cudaPointerGetAttributes(&pa, src) == cudaSuccess) {
                if (pa.type == cudaMemoryTypeHost) max_ctas = kXfHostCtas;
            } else {
                (void)cudaGetLastError();
            }
        }
        rc = dst_dtype == TB_F32 ? launch_stream<float>(p, dev, s, fr, fw, max_ctas, pdl, &taken)
                                 : launch_stream<uint16_t>(p, dev, s, fr, fw, max_ctas, pdl, &taken);
        if (rc) return rc;
        if (taken) return TB_OK;
    }
    const bool pdl_generic = !(flags & TB_FLAG_NO_PDL);
    if (dst_dtype == TB_F32) return launch_generic<float>(p, s, pdl_generic);
    return launch_generic<uint16_t>(p, s, pdl_generic);
}

extern "C" int tb_transform_u8_nhwc_to_nchw(const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride,
                                            int64_t src_row_pitch, void* dst, int dst_dtype, int out_h, int out_w,
                                            const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
                                            int pad_y, int pad_x, int pad_mode, const void* lut, int flags,
                                            tb_stream_t stream) {
    return tb_transform_impl(src, nullptr, B, H, W, C, src_batch_stride, src_row_pitch, dst, dst_dtype, out_h, out_w,
                             crop_top, crop_left, flip, pad_y, pad_x, pad_mode, lut, flags, stream);
}

// Same transform over records that lie at arbitrary offsets from one base (an LMDB map registered
// with tb_lmdb_cuda_register): image b starts at src + offsets[b].  Offsets must be multiples of 16
// for the streaming path (the caller guarantees it; LMDB overflow values are).
extern "C" int tb_transform_u8_records_to_nchw(const uint8_t* src, const int64_t* offsets, int B, int H, int W, int C,
                                               int64_t src_row_pitch, void* dst, int dst_dtype, int out_h, int out_w,
                                               const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
                                               int pad_y, int pad_x, int pad_mode, const void* lut, int flags,
                                               tb_stream_t stream) {
    TB_REQUIRE(offsets != nullptr || B == 0, "tb_transform_u8_records: offsets is NULL");
    return tb_transform_impl(src, offsets, B, H, W, C, 0, src_row_pitch, dst, dst_dtype, out_h, out_w, crop_top, crop_left,
                             flip, pad_y, pad_x, pad_mode, lut, flags, stream);
}
