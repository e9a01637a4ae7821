// One pass over a logits matrix: top-1 / top-k correct counts + confusion bins (sm_100a).
//
// Replaces the reference's `(logits.argmax(dim) == labels).sum()` (torchbooster/metrics.py:27:
// argmax, eq, sum = three eager kernels plus a host sync at examples/img_cls/resnet/resnet.py:67)
// and adds the top-k / confusion accumulators of the north star (rank rule, DESIGN.md "metrics").
//
// One warp per row.  Every element is mapped to a signed 32-bit key that orders like the float
// with NaN maximal and -0 == +0; the label's rank is counted and the argmax tracked in the same
// sweep; three redux.sync per row finish it.  Row sweeps are 16-byte loads; rows that are not
// 16-byte aligned get scalar head / tail elements, never an out-of-row vector load.
#include "tb_common.cuh"

namespace {

constexpr int kMxThreads = 256;
constexpr int kMxWarps = kMxThreads / 32;

// Order-preserving key.  `x + 0.0f` turns -0 into +0 and any NaN into the canonical
// 0x7fffffff (PTX: f32 arithmetic returns the canonical NaN), which maps to INT_MAX.
__device__ __forceinline__ int tb_key(float x) {
    const int s = __float_as_int(x + 0.0f);
    return s ^ ((s >> 31) & 0x7fffffff);
}

template <int DT>
struct Elem;
template <>
struct Elem<TB_F32> {
    static constexpr int VE = 4;
    using type = float;
    __device__ static __forceinline__ float get(const void* p, int64_t i) { return ((const float*)p)[i]; }
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    }
};
template <>
struct Elem<TB_BF16> {
    static constexpr int VE = 8;
    using type = uint16_t;
    __device__ static __forceinline__ float get(const void* p, int64_t i) {
        return __uint_as_float((uint32_t)((const uint16_t*)p)[i] << 16);
    }
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
};
template <>
struct Elem<TB_F16> {
    static constexpr int VE = 8;
    using type = uint16_t;
    __device__ static __forceinline__ float get(const void* p, int64_t i) { return __half2float(((const __half*)p)[i]); }
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = t.x;
            f[2 * i + 1] = t.y;
        }
    }
};

// ---- per-row accumulator ---------------------------------------------------------------------
// The sweep keeps per lane: `cnt` (how many elements outrank the label), and the key and position
// of the best vector seen (first one wins); the exact argmax index is resolved once per row by
// re-reading that one vector.
//
// Rank counting runs on the float pipe.  For a label logit ly that is not NaN,
//   j <  y counts when x_j >= ly  <=>  x_j > pred(ly)   (pred = next float below; NaN x_j counts)
//   j >= y counts when x_j >  ly                         (NaN x_j counts)
// so one "greater than or unordered" compare against a per-vector threshold does both; the one
// vector that straddles y selects per element.  A NaN label logit takes the key-based route.
__device__ __forceinline__ int tb_gtu(float x, float thr) {   // -1 when x > thr or unordered, else 0
    int r;
    asm("set.gtu.s32.f32 %0, %1, %2;" : "=r"(r) : "f"(x), "f"(thr));
    return r;
}
__device__ __forceinline__ float tb_max_nan(float a, float b) {   // NaN-propagating max
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

struct RowAcc {
    int bestk, bj0, bn;     // key of the best vector, its first element index and length
    int cnt;                // minus the number of elements that outrank the label
    int y, ky;
    float lo, hi;           // thresholds for j < y and j >= y
    bool ynan;

    __device__ __forceinline__ void init(float ly, int label) {
        y = label;
        ky = tb_key(ly);
        ynan = ly != ly;
        const float c = ly + 0.0f;
        const uint32_t b = __float_as_uint(c);
        hi = c;
        if (b == 0xff800000u) lo = __uint_as_float(0x7fffffffu);            // >= -inf: everything counts
        else if (b == 0u) lo = __uint_as_float(0x80000001u);                // below zero: -denorm_min
        else lo = __uint_as_float((b & 0x80000000u) ? b + 1u : b - 1u);
        bestk = INT_MIN; bj0 = INT_MAX; bn = 0; cnt = 0;
    }
    // scalar element (row head / tail, NaN-label rows)
    __device__ __forceinline__ void one(float x, int j) {
        const int k = tb_key(x);
        cnt -= (k > ky - (j < y ? 1 : 0)) ? 1 : 0;
        if (k > bestk) { bestk = k; bj0 = j; bn = 1; }
    }
    template <int VE>
    __device__ __forceinline__ void vec(const float* f, int j0) {
        float m = tb_max_nan(f[0], f[1]);
#pragma unroll
        for (int e = 2; e < VE; ++e) m = tb_max_nan(m, f[e]);
        const int k = tb_key(m);
        if (k > bestk) { bestk = k; bj0 = j0; bn = VE; }
        const int d = y - j0;
        if ((unsigned)d < (unsigned)VE) {
#pragma unroll
            for (int e = 0; e < VE; ++e) cnt += tb_gtu(f[e], e < d ? lo : hi);
        } else {
            const float thr = d > 0 ? lo : hi;
#pragma unroll
            for (int e = 0; e < VE; ++e) cnt += tb_gtu(f[e], thr);
        }
    }
};

template <int DT>
__global__ void __launch_bounds__(kMxThreads) tb_metrics_kernel(const void* __restrict__ logits, int64_t N, int C,
                                                                int64_t row_stride, const int64_t* __restrict__ labels,
                                                                int k, unsigned long long* __restrict__ counters,
                                                                unsigned long long* __restrict__ confusion) {
    using E = Elem<DT>;
    constexpr int VE = E::VE;
    constexpr int ES = 16 / VE;
    constexpr int UN = 32 / VE;      // 16-byte loads per lane per batch: 8 for fp32, 4 for 16-bit types
    __shared__ unsigned int s_cnt[3];
    if (threadIdx.x < 3) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    tb_launch_dependents();
    tb_grid_dependency_wait();      // first global access below (programmatic dependent launch)

    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * kMxWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kMxWarps;
    unsigned int c1 = 0, ck = 0, inv = 0;

    for (int64_t n = warp0; n < N; n += nwarps) {
        const long long y64 = labels[n];
        if (y64 < 0 || y64 >= C) { inv += 1; continue; }
        const typename E::type* row = (const typename E::type*)logits + n * row_stride;
        RowAcc a;
        a.init(E::get(row, (int)y64), (int)y64);

        const int mis = (int)(((uintptr_t)row & 15) / ES);          // elements past the previous 16-byte boundary
        const int jf = mis ? min(VE - mis, C) : 0;                  // head elements before the first aligned vector
        const int nfull = (C - jf) / VE;
        const int jt = jf + nfull * VE;                             // first tail element
        const uint4* vp = reinterpret_cast<const uint4*>(row + jf);

        if (a.ynan) {
            // rare: the label's own logit is NaN -- everything through the key route
            for (int j = lane; j < C; j += 32) a.one(E::get(row, j), j);
        } else {
            if (lane < jf) a.one(E::get(row, lane), lane);
            // UN vectors per lane in flight at once (a 1000-class fp32 row is 250 vectors: one batch)
            for (int v = lane; v < nfull; v += 32 * UN) {
                uint4 r[UN];
#pragma unroll
                for (int i = 0; i < UN; ++i)
                    if (v + 32 * i < nfull) r[i] = __ldcs(vp + v + 32 * i);
#pragma unroll
                for (int i = 0; i < UN; ++i) {
                    if (v + 32 * i < nfull) {
                        float f[VE];
                        E::unpack(r[i], f);
                        a.template vec<VE>(f, jf + (v + 32 * i) * VE);
                    }
                }
            }
            if (jt + lane < C) a.one(E::get(row, jt + lane), jt + lane);   // tail: fewer than VE <= 32 elements
        }

        const int mx = __reduce_max_sync(0xffffffffu, a.bestk);
        int cand = INT_MAX;
        if (a.bestk == mx) {
            // first element of this lane's best vector that attains the maximum
            for (int e = a.bn - 1; e >= 0; --e)
                if (tb_key(E::get(row, a.bj0 + e)) == mx) cand = a.bj0 + e;
        }
        const int pred = __reduce_min_sync(0xffffffffu, cand);
        const int rank = -__reduce_add_sync(0xffffffffu, a.cnt);
        if (lane == 0) {
            c1 += (rank < 1) ? 1u : 0u;
            ck += (rank < k) ? 1u : 0u;
            if (confusion) atomicAdd(&confusion[(int64_t)a.y * C + pred], 1ULL);
        }
    }
    if (lane == 0) {
        if (c1) atomicAdd(&s_cnt[0], c1);
        if (ck) atomicAdd(&s_cnt[1], ck);
        if (inv) atomicAdd(&s_cnt[2], inv);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) atomicAdd(&counters[0], (unsigned long long)N);
        if (s_cnt[0]) atomicAdd(&counters[1], (unsigned long long)s_cnt[0]);
        if (s_cnt[1]) atomicAdd(&counters[2], (unsigned long long)s_cnt[1]);
        if (s_cnt[2]) atomicAdd(&counters[3], (unsigned long long)s_cnt[2]);
    }
}

// Deterministic float64 sum of a short vector (per-step losses) into a device-resident state.
template <int DT>
__global__ void __launch_bounds__(256) tb_running_sum_kernel(const void* __restrict__ x, int64_t n,
                                                             double* __restrict__ state) {
    __shared__ double s[256];
    double acc = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += 256) acc += (double)Elem<DT>::get(x, i);
    s[threadIdx.x] = acc;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        state[0] += s[0];
        state[1] += (double)n;
    }
}

}  // namespace

extern "C" int tb_metrics_update(const void* logits, int logits_dtype, int64_t N, int C, int64_t row_stride,
                                 const int64_t* labels, int k, int64_t* counters, int64_t* confusion,
                                 tb_stream_t stream) {
    TB_REQUIRE(N >= 0, "tb_metrics_update: negative N");
    TB_REQUIRE(C >= 1, "tb_metrics_update: C=%d must be >= 1", C);
    TB_REQUIRE(k >= 1, "tb_metrics_update: k=%d must be >= 1", k);
    TB_REQUIRE(row_stride >= C, "tb_metrics_update: row_stride %lld < C", (long long)row_stride);
    const int es = tb_dtype_size(logits_dtype);
    TB_REQUIRE(es != 0, "tb_metrics_update: unknown logits_dtype %d", logits_dtype);
    TB_REQUIRE(counters != nullptr, "tb_metrics_update: counters is NULL");
    if (N == 0) return TB_OK;
    TB_REQUIRE(logits && labels, "tb_metrics_update: NULL logits/labels");
    TB_REQUIRE(((uintptr_t)logits % es) == 0, "tb_metrics_update: logits misaligned for its dtype");
    // one sample per warp per block-round keeps the 32-bit block counters far from overflow
    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;
    int64_t blocks = (N + kMxWarps - 1) / kMxWarps;
    const int64_t cap = (int64_t)dev.sm_count * 8;   // 8 x 256 threads = full occupancy
    if (blocks > cap) blocks = cap;
    TB_REQUIRE(N / (blocks * kMxWarps) < ((int64_t)1 << 28), "tb_metrics_update: N too large for one launch");
    auto* cnt = reinterpret_cast<unsigned long long*>(counters);
    auto* cm = reinterpret_cast<unsigned long long*>(confusion);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(kMxThreads);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (logits_dtype == TB_F32)
        TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_kernel<TB_F32>, logits, N, C, row_stride, labels, k, cnt, cm));
    else if (logits_dtype == TB_BF16)
        TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_kernel<TB_BF16>, logits, N, C, row_stride, labels, k, cnt, cm));
    else
        TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_kernel<TB_F16>, logits, N, C, row_stride, labels, k, cnt, cm));
    TB_LAUNCH_CHECK("tb_metrics_kernel");
    return TB_OK;
}

extern "C" int tb_running_sum_update(const void* x, int dtype, int64_t n, double* state, tb_stream_t stream) {
    TB_REQUIRE(n >= 0, "tb_running_sum_update: negative n");
    TB_REQUIRE(tb_dtype_size(dtype) != 0, "tb_running_sum_update: unknown dtype %d", dtype);
    TB_REQUIRE(state != nullptr, "tb_running_sum_update: state is NULL");
    if (n == 0) return TB_OK;
    TB_REQUIRE(x != nullptr, "tb_running_sum_update: x is NULL");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == TB_F32) tb_running_sum_kernel<TB_F32><<<1, 256, 0, s>>>(x, n, state);
    else if (dtype == TB_BF16) tb_running_sum_kernel<TB_BF16><<<1, 256, 0, s>>>(x, n, state);
    else tb_running_sum_kernel<TB_F16><<<1, 256, 0, s>>>(x, n, state);
    TB_LAUNCH_CHECK("tb_running_sum_kernel");
    return TB_OK;
}
