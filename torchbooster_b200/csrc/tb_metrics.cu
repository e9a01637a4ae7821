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
    __device__ static __forceinline__ float to_float(float v) { return v; }
    __device__ static __forceinline__ float get(const void* p, int64_t i) { return ((const float*)p)[i]; }
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    }
};
template <>
struct Elem<TB_BF16> {
    static constexpr int VE = 8;
    using type = uint16_t;
    __device__ static __forceinline__ float to_float(uint16_t v) { return __uint_as_float((uint32_t)v << 16); }
    __device__ static __forceinline__ float get(const void* p, int64_t i) { return to_float(((const uint16_t*)p)[i]); }
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
    __device__ static __forceinline__ float to_float(uint16_t v) { return __half2float(__ushort_as_half(v)); }
    __device__ static __forceinline__ float get(const void* p, int64_t i) { return to_float(((const uint16_t*)p)[i]); }
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
// The sweep keeps per lane: `cnt` (how many elements outrank the label, scaled by kScale) and the
// value, key and position of the best 16-byte vector seen (first one wins); the exact argmax
// index is resolved once per row by re-reading that one vector.
//
// Rank counting needs no integer keys.  For a label logit ly that is not NaN,
//   j <  y counts when x_j >= ly  <=>  x_j > pred(ly)   (pred = next value below; NaN x_j counts)
//   j >= y counts when x_j >  ly                         (NaN x_j counts)
// so one "greater than or unordered" compare against a per-vector threshold does both; the one
// vector that straddles y selects per element.  16-bit logits are compared two at a time
// (set.gtu.{bf16x2,f16x2}) and the eight half-word masks of a vector are folded into one word
// whose popcount is 4x the number of hits.  A NaN label logit takes the key-based route.
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
__device__ __forceinline__ uint32_t tb_bitsel(uint32_t a, uint32_t b, uint32_t k) { return (a & k) | (b & ~k); }

template <int DT>
struct Packed {   // fp32: not packed
    static constexpr bool kOn = false;
    static constexpr int kScale = 1;
};
template <>
struct Packed<TB_BF16> {
    static constexpr bool kOn = true;
    static constexpr int kScale = 4;
    static constexpr uint32_t kNegInf = 0xff80u;
    __device__ static __forceinline__ uint32_t gtu2(uint32_t a, uint32_t b) {   // 0xffff per half-word that is > or unordered
        uint32_t r;
        asm("set.gtu.u32.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }
    __device__ static __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
        uint32_t r;
        asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }
    __device__ static __forceinline__ uint32_t eq2(uint32_t a, uint32_t b) {    // 0xffff per half-word that compares equal (-0 == +0)
        uint32_t r;
        asm("set.eq.u32.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }
    __device__ static __forceinline__ uint32_t bits(float m) { return __float_as_uint(m) >> 16; }      // exact: m came from a bf16
    __device__ static __forceinline__ float lo(uint32_t w) { return __uint_as_float(w << 16); }
    __device__ static __forceinline__ float hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
};
template <>
struct Packed<TB_F16> {
    static constexpr bool kOn = true;
    static constexpr int kScale = 4;
    static constexpr uint32_t kNegInf = 0xfc00u;
    __device__ static __forceinline__ uint32_t gtu2(uint32_t a, uint32_t b) {
        uint32_t r;
        asm("set.gtu.u32.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }
    __device__ static __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
        uint32_t r;
        asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }
    __device__ static __forceinline__ uint32_t eq2(uint32_t a, uint32_t b) {
        uint32_t r;
        asm("set.eq.u32.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }
    __device__ static __forceinline__ uint32_t bits(float m) { return (uint32_t)__half_as_ushort(__float2half_rn(m)); }   // exact
    __device__ static __forceinline__ float lo(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)w)); }
    __device__ static __forceinline__ float hi(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w >> 16))); }
};

template <int DT>
struct RowAcc {
    using E = Elem<DT>;
    using P = Packed<DT>;
    static constexpr int VE = E::VE;
    int bestk, bj0, bn;     // key of the best vector, its first element index and length
    float bm;               // value behind bestk (NaN until the first update)
    int cnt;                // kScale x the number of elements that outrank the label
    int y, ky;
    float lo, hi;           // thresholds for j < y and j >= y
    uint32_t lo2, hi2;      // the same, as packed pairs (16-bit logits)
    bool ynan;

    __device__ __forceinline__ void init(typename E::type raw, int label) {
        y = label;
        const float ly = E::to_float(raw);
        ky = tb_key(ly);
        ynan = ly != ly;
        bestk = INT_MIN; bj0 = INT_MAX; bn = 0; cnt = 0;
        bm = __uint_as_float(0x7fffffffu);
        lo2 = hi2 = 0;
        if constexpr (P::kOn) {
            // thresholds in the logits' own 16-bit format (exact in fp32 as well)
            uint32_t c = (uint32_t)raw;
            if (c == 0x8000u) c = 0;                                   // -0 == +0
            uint32_t l;
            if (c == P::kNegInf) l = 0x7fffu;                          // >= -inf: everything counts
            else if (c == 0u) l = 0x8001u;                             // just below zero
            else l = (c & 0x8000u) ? c + 1u : c - 1u;
            hi2 = c * 0x10001u;
            lo2 = l * 0x10001u;
            hi = P::lo(hi2);
            lo = P::lo(lo2);
        } else {
            const float c = ly + 0.0f;
            const uint32_t b = __float_as_uint(c);
            hi = c;
            if (b == 0xff800000u) lo = __uint_as_float(0x7fffffffu);
            else if (b == 0u) lo = __uint_as_float(0x80000001u);
            else lo = __uint_as_float((b & 0x80000000u) ? b + 1u : b - 1u);
        }
    }
    // scalar element (row head / tail, NaN-label rows)
    __device__ __forceinline__ void one(float x, int j) {
        const int k = tb_key(x);
        cnt += (k > ky - (j < y ? 1 : 0)) ? P::kScale : 0;
        if (k > bestk) { bestk = k; bm = x; bj0 = j; bn = 1; }
    }
    __device__ __forceinline__ void candidate(float m, int j0) {
        if (!(m <= bm)) {                     // greater or unordered: rare after the first vectors
            const int k = tb_key(m);
            if (k > bestk) { bestk = k; bm = m; bj0 = j0; bn = VE; }
        }
    }
    // NaN-propagating maximum of one 16-byte vector
    __device__ __forceinline__ float vec_max(const uint4& r) const {
        if constexpr (P::kOn) {
            const uint32_t m = P::max2(P::max2(r.x, r.y), P::max2(r.z, r.w));
            return tb_max_nan(P::lo(m), P::hi(m));
        } else {
            float f[VE];
            E::unpack(r, f);
            float m = tb_max_nan(f[0], f[1]);
#pragma unroll
            for (int e = 2; e < VE; ++e) m = tb_max_nan(m, f[e]);
            return m;
        }
    }
    // elements of one vector that outrank the label
    __device__ __forceinline__ void vec_count(const uint4& r, int j0) {
        const int d = y - j0;
        if ((unsigned)d < (unsigned)VE) {        // the one vector that straddles the label
            float f[VE];
            E::unpack(r, f);
#pragma unroll
            for (int e = 0; e < VE; ++e) cnt -= P::kScale * tb_gtu(f[e], e < d ? lo : hi);
        } else if constexpr (P::kOn) {
            const uint32_t thr = d > 0 ? lo2 : hi2;
            // nibble i of each half-word <- hit mask of word i; popcount = 4 x hits
            const uint32_t t = tb_bitsel(tb_bitsel(P::gtu2(r.x, thr), P::gtu2(r.y, thr), 0x000f000fu),
                                         tb_bitsel(P::gtu2(r.z, thr), P::gtu2(r.w, thr), 0x0f000f00u), 0x00ff00ffu);
            cnt += __popc(t);
        } else {
            float f[VE];
            E::unpack(r, f);
            const float thr = d > 0 ? lo : hi;
#pragma unroll
            for (int e = 0; e < VE; ++e) cnt -= tb_gtu(f[e], thr);
        }
    }
    // one batch of UN vectors (vector i holds elements j0 + i * stride ...): count every vector, but
    // rank only the batch maximum against the best so far
    template <int UN>
    __device__ __forceinline__ void batch(const uint4* r, int v, int nfull, int jf) {
        const float ninf = __uint_as_float(0xff800000u);
        float m[UN];
#pragma unroll
        for (int i = 0; i < UN; ++i) {
            const bool ok = v + 32 * i < nfull;
            m[i] = ok ? vec_max(r[i]) : ninf;
            if (ok) vec_count(r[i], jf + (v + 32 * i) * VE);
        }
        float M = m[0];
#pragma unroll
        for (int i = 1; i < UN; ++i) M = tb_max_nan(M, m[i]);
        if (v < nfull && !(M <= bm)) {       // (vector 0 invalid = no vector of this lane is valid)
            const bool nanmax = M != M;
            int pick = 0;
#pragma unroll
            for (int i = UN - 1; i >= 0; --i)
                if (nanmax ? (m[i] != m[i]) : (m[i] >= M)) pick = i;
            candidate(M, jf + (v + 32 * pick) * VE);
        }
    }
    // first index attaining the row maximum inside this lane's best vector
    __device__ __forceinline__ int resolve(const typename E::type* row) const {
        if (bn == 1) return bj0;
        const uint4 r = *reinterpret_cast<const uint4*>(row + bj0);
        float f[VE];
        E::unpack(r, f);
        const bool nanmax = bm != bm;
        int e0 = 0;
#pragma unroll
        for (int e = VE - 1; e >= 0; --e)
            if (nanmax ? (f[e] != f[e]) : (f[e] >= bm)) e0 = e;
        return bj0 + e0;
    }
};

// ALIGNED: every row starts on a 16-byte boundary (base and row stride are multiples of 16
// bytes), so the head / body / tail split is the same for all rows.
// SCM: the confusion bins are privatised per CTA in shared memory (uint32 [C * C], C * C * 4 <= kScmMaxBytes): a sample is
// one shared-memory atomic, and each CTA adds its non-zero bins to the global matrix once at the end.  With few classes
// (the reference's workloads have 10, examples/img_cls/resnet/resnet.yml) every sample of a global-atomic version
// lands on one of C * C hot addresses.
constexpr int kScmMaxBytes = 48 * 1024;
template <int DT, bool ALIGNED, bool SCM>
__global__ void __launch_bounds__(kMxThreads, 4) tb_metrics_kernel(const void* __restrict__ logits, int64_t N, int C,
                                                                int64_t row_stride, const int64_t* __restrict__ labels,
                                                                int k, unsigned long long* __restrict__ counters,
                                                                unsigned long long* __restrict__ confusion) {
    using E = Elem<DT>;
    constexpr int VE = E::VE;
    constexpr int ES = 16 / VE;
    constexpr int UN = 32 / VE;      // 16-byte loads per lane per batch: 8 for fp32, 4 for 16-bit types
    __shared__ unsigned int s_cnt[3];
    extern __shared__ unsigned int s_cm[];        // SCM only: C * C bins
    if (threadIdx.x < 3) s_cnt[threadIdx.x] = 0;
    if (SCM) {
        for (int i = threadIdx.x; i < C * C; i += kMxThreads) s_cm[i] = 0;
    }
    __syncthreads();
    tb_launch_dependents();
    tb_grid_dependency_wait();      // first global access below (programmatic dependent launch)

    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * kMxWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kMxWarps;
    unsigned int c1 = 0, ck = 0, inv = 0;

    // One DRAM round trip per row: the next row's label is fetched a row ahead, and the row's
    // vectors, its head / tail elements and the label's own logit are all issued before any of
    // them is used.
    long long y64 = warp0 < N ? labels[warp0] : -1;
    for (int64_t n = warp0; n < N; n += nwarps) {
        const long long ycur = y64;
        if (n + nwarps < N) y64 = labels[n + nwarps];
        if (ycur < 0 || ycur >= C) { inv += 1; continue; }
        const int y = (int)ycur;
        const typename E::type* row = (const typename E::type*)logits + n * row_stride;

        const int mis = ALIGNED ? 0 : (int)(((uintptr_t)row & 15) / ES);   // elements past the previous 16-byte boundary
        const int jf = mis ? min(VE - mis, C) : 0;                  // head elements before the first aligned vector
        const int nfull = (C - jf) / VE;
        const int jt = jf + nfull * VE;                             // first tail element
        const uint4* vp = reinterpret_cast<const uint4*>(row + jf);

        // UN vectors per lane in flight at once (a 1000-class fp32 row is 250 vectors: one batch)
        uint4 r[UN];
#pragma unroll
        for (int i = 0; i < UN; ++i)
            if (lane + 32 * i < nfull) r[i] = __ldcs(vp + lane + 32 * i);
        const typename E::type raw_y = row[y];
        typename E::type raw_head = raw_y, raw_tail = raw_y;
        if (lane < jf) raw_head = row[lane];
        if (jt + lane < C) raw_tail = row[jt + lane];                 // tail: fewer than VE <= 32 elements

        RowAcc<DT> a;
        a.init(raw_y, y);
        if (a.ynan) {
            // rare: the label's own logit is NaN -- everything through the key route
            for (int j = lane; j < C; j += 32) a.one(E::get(row, j), j);
        } else {
            if (lane < jf) a.one(E::to_float(raw_head), lane);
            a.template batch<UN>(r, lane, nfull, jf);
            for (int v = lane + 32 * UN; v < nfull; v += 32 * UN) {   // rows longer than one batch
#pragma unroll
                for (int i = 0; i < UN; ++i)
                    if (v + 32 * i < nfull) r[i] = __ldcs(vp + v + 32 * i);
                a.template batch<UN>(r, v, nfull, jf);
            }
            if (jt + lane < C) a.one(E::to_float(raw_tail), jt + lane);
        }

        const int mx = __reduce_max_sync(0xffffffffu, a.bestk);
        const int cand = a.bestk == mx ? a.resolve(row) : INT_MAX;
        const int pred = __reduce_min_sync(0xffffffffu, cand);
        const int rank = __reduce_add_sync(0xffffffffu, a.cnt) / Packed<DT>::kScale;
        if (lane == 0) {
            c1 += (rank < 1) ? 1u : 0u;
            ck += (rank < k) ? 1u : 0u;
            if (SCM) atomicAdd(&s_cm[a.y * C + pred], 1u);
            else if (confusion) atomicAdd(&confusion[(int64_t)a.y * C + pred], 1ULL);
        }
    }
    if (lane == 0) {
        if (c1) atomicAdd(&s_cnt[0], c1);
        if (ck) atomicAdd(&s_cnt[1], ck);
        if (inv) atomicAdd(&s_cnt[2], inv);
    }
    __syncthreads();
    if (SCM) {
        for (int i = threadIdx.x; i < C * C; i += kMxThreads) {
            const unsigned int v = s_cm[i];
            if (v) atomicAdd(&confusion[i], (unsigned long long)v);
        }
    }
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) atomicAdd(&counters[0], (unsigned long long)N);
        if (s_cnt[0]) atomicAdd(&counters[1], (unsigned long long)s_cnt[0]);
        if (s_cnt[1]) atomicAdd(&counters[2], (unsigned long long)s_cnt[1]);
        if (s_cnt[2]) atomicAdd(&counters[3], (unsigned long long)s_cnt[2]);
    }
}

// ---- lean variant for rows that are ONE batch: every row 16-byte aligned, C a multiple of the vector width and at most
// 32 * UN vectors per row (C <= 1024 for every dtype: ImageNet's 1000 classes).  Same rank rule and argmax, a third of the
// instructions per row: the general kernel above spends ~430 (bf16) / ~550 (fp32) warp instructions on a 1000-class row,
// most of them on head / tail / multi-batch bookkeeping that such rows never need, and was issue-bound (0.52 / 0.78 of
// the HBM peak).  Differences in method, not in result:
//   * lanes past the end of the row hold -inf vectors: they neither count (x > thr is false for the thresholds such
//     vectors get) nor win the maximum, so nothing below tests "is this vector valid";
//   * every vector is counted against ONE threshold -- pred(l_y) when it lies wholly before the label, l_y otherwise --
//     and the one vector that straddles the label adds the ties among its elements before the label separately;
//   * the argmax is the warp maximum of the per-lane keys, then the smallest index among the lanes that hold it,
//     found in registers (no second read of the row).
__device__ __forceinline__ float tb_max3_nan(float a, float b, float c) {
    float r;
    asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ uint4 tb_ldcs_or(const uint4* p, bool take, uint32_t fill) {
    uint4 v = make_uint4(fill, fill, fill, fill);
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\t@q ld.global.cs.v4.u32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w) : "l"(p), "r"((int)take));
    return v;
}

// volatile: the row's loads stay together at the top of the iteration, all in flight before the first use (left to the
// scheduler they sink towards their uses and a row costs several DRAM round trips instead of one)
__device__ __forceinline__ uint4 tb_ldcs_now(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.cs.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

template <int DT, bool SCM>
__global__ void __launch_bounds__(kMxThreads, 4) tb_metrics_row_kernel(const void* __restrict__ logits, int64_t N, int C,
                                                                    int64_t row_stride, const int64_t* __restrict__ labels,
                                                                    int k, unsigned long long* __restrict__ counters,
                                                                    unsigned long long* __restrict__ confusion) {
    using E = Elem<DT>;
    using P = Packed<DT>;
    constexpr int VE = E::VE;
    constexpr int UN = 32 / VE;      // vectors per lane: 8 for fp32, 4 for 16-bit types (1024 elements per row at most)
    constexpr uint32_t kFill = DT == TB_F32 ? 0xff800000u : (DT == TB_BF16 ? 0xff80ff80u : 0xfc00fc00u);     // -inf
    __shared__ unsigned int s_cnt[3];
    extern __shared__ unsigned int s_cm[];        // SCM only: C * C bins
    if (threadIdx.x < 3) s_cnt[threadIdx.x] = 0;
    if (SCM) {
        for (int i = threadIdx.x; i < C * C; i += kMxThreads) s_cm[i] = 0;
    }
    __syncthreads();
    tb_launch_dependents();
    tb_grid_dependency_wait();

    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * kMxWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kMxWarps;
    const int nv = C / VE;
    unsigned int c1 = 0, ck = 0, inv = 0;
    bool take[UN], whole[UN];       // this lane has vector i; every lane has vector i (no -inf fill needed)
#pragma unroll
    for (int i = 0; i < UN; ++i) { take[i] = lane + 32 * i < nv; whole[i] = 32 * (i + 1) <= nv; }

    long long y64 = warp0 < N ? labels[warp0] : -1;
    const typename E::type* row = (const typename E::type*)logits + warp0 * row_stride;
    const int64_t row_step = nwarps * row_stride;
    for (int64_t n = warp0; n < N; n += nwarps, row += row_step) {
        const long long ycur = y64;
        if (n + nwarps < N) y64 = labels[n + nwarps];
        if (ycur < 0 || ycur >= C) { inv += 1; continue; }
        const int y = (int)ycur;
        const uint4* vp = reinterpret_cast<const uint4*>(row) + lane;
        uint4 r[UN];
#pragma unroll
        for (int i = 0; i < UN; ++i) {
            if (whole[i]) r[i] = tb_ldcs_now(vp + 32 * i);
            else r[i] = tb_ldcs_or(vp + 32 * i, take[i], kFill);
        }
        const typename E::type raw_y = row[y];
        // the elements that share the label's 16-byte vector and lie before it, one per lane (ties among them count)
        const int d0 = y & (VE - 1);
        const typename E::type raw_t = row[y - d0 + min(lane, VE - 1)];
        const float ly = E::to_float(raw_y);
        if (ly != ly) {
            // rare: the label's own logit is NaN -- the key route of the general kernel, element by element
            RowAcc<DT> a;
            a.init(raw_y, y);
            for (int j = lane; j < C; j += 32) a.one(E::get(row, j), j);
            const int mx = __reduce_max_sync(0xffffffffu, a.bestk);
            const int pred = __reduce_min_sync(0xffffffffu, a.bestk == mx ? a.bj0 : INT_MAX);
            const int rank = __reduce_add_sync(0xffffffffu, a.cnt) / P::kScale;
            if (lane == 0) {
                c1 += (rank < 1) ? 1u : 0u;
                ck += (rank < k) ? 1u : 0u;
                if (SCM) atomicAdd(&s_cm[y * C + pred], 1u);
                else if (confusion) atomicAdd(&confusion[(int64_t)y * C + pred], 1ULL);
            }
            continue;
        }

        // ---- thresholds: x > hi  <=>  x outranks the label from behind it; x > lo  <=>  x >= l_y (from before it).
        // A vector wholly before the label is counted against lo, every other one against hi; the ties before the label
        // inside its own vector are the (lane < d0) term.
        int cnt = (lane < d0 && E::to_float(raw_t) == ly) ? P::kScale : 0;
        float m[UN];
        const int dl = y - lane * VE;                       // label position relative to this lane's first vector
        if constexpr (P::kOn) {
            uint32_t c = (uint32_t)raw_y;
            if (c == 0x8000u) c = 0;                                   // -0 == +0
            uint32_t l;
            if (c == P::kNegInf) l = 0x7fffu;                          // >= -inf: everything counts (NaN threshold: unordered)
            else if (c == 0u) l = 0x8001u;                             // just below zero
            else l = (c & 0x8000u) ? c + 1u : c - 1u;
            const uint32_t hi2 = c * 0x10001u, lo2 = l * 0x10001u;
#pragma unroll
            for (int i = 0; i < UN; ++i) {
                const int d = dl - 32 * i * VE;                        // >= VE: the whole vector lies before the label
                const uint32_t thr = d >= VE ? lo2 : hi2;
                const uint32_t t = tb_bitsel(tb_bitsel(P::gtu2(r[i].x, thr), P::gtu2(r[i].y, thr), 0x000f000fu),
                                             tb_bitsel(P::gtu2(r[i].z, thr), P::gtu2(r[i].w, thr), 0x0f000f00u), 0x00ff00ffu);
                cnt += __popc(t);
                const uint32_t mm = P::max2(P::max2(r[i].x, r[i].y), P::max2(r[i].z, r[i].w));
                m[i] = tb_max_nan(P::lo(mm), P::hi(mm));
            }
        } else {
            const float cc = ly + 0.0f;
            const uint32_t b = __float_as_uint(cc);
            const float hi = cc;
            float lo;
            if (b == 0xff800000u) lo = __uint_as_float(0x7fffffffu);
            else if (b == 0u) lo = __uint_as_float(0x80000001u);
            else lo = __uint_as_float((b & 0x80000000u) ? b + 1u : b - 1u);
#pragma unroll
            for (int i = 0; i < UN; ++i) {
                const int d = dl - 32 * i * VE;
                const float thr = d >= VE ? lo : hi;
                const float f0 = __uint_as_float(r[i].x), f1 = __uint_as_float(r[i].y), f2 = __uint_as_float(r[i].z),
                            f3 = __uint_as_float(r[i].w);
                cnt = cnt - tb_gtu(f0, thr) - tb_gtu(f1, thr);
                cnt = cnt - tb_gtu(f2, thr) - tb_gtu(f3, thr);
                m[i] = tb_max_nan(tb_max3_nan(f0, f1, f2), f3);
            }
        }
        // ---- argmax: warp maximum of the keys, then the first index among the lanes that hold it
        float M = m[0];
#pragma unroll
        for (int i = 1; i < UN; ++i) M = tb_max_nan(M, m[i]);
        const int key = tb_key(M);
        const int mx = __reduce_max_sync(0xffffffffu, key);
        int cand = INT_MAX;
        if (key == mx) {
            if (M != M) {
                // a NaN is the maximum (rare): first NaN of this lane
                int pick = 0;
#pragma unroll
                for (int i = UN - 1; i >= 0; --i)
                    if (m[i] != m[i]) pick = i;
                if (lane + 32 * pick < nv) {
                    const uint4 rv = __ldg(vp + 32 * pick);
                    float f[VE];
                    E::unpack(rv, f);
                    int e0 = 0;
#pragma unroll
                    for (int e = VE - 1; e >= 0; --e)
                        if (f[e] != f[e]) e0 = e;
                    cand = (lane + 32 * pick) * VE + e0;
                }
            } else {
                int pick = 0;
#pragma unroll
                for (int i = UN - 1; i >= 0; --i)
                    if (m[i] >= M) pick = i;
                // The winning vector is read again (one 16-byte load by one lane, an L2 hit) instead of being kept:
                // holding all UN vectors for this moment cost the fp32 kernel its registers -- the compiler reused load
                // registers and a row took three DRAM round trips instead of one.
                if (lane + 32 * pick < nv) {
                const uint4 rv = __ldg(vp + 32 * pick);
                int e0;
                if constexpr (P::kOn) {
                    // packed equality against the maximum; nibble w of the low / high half <- word w's low / high element
                    const uint32_t mp = P::bits(M) * 0x10001u;
                    const uint32_t t = tb_bitsel(tb_bitsel(P::eq2(rv.x, mp), P::eq2(rv.y, mp), 0x000f000fu),
                                                 tb_bitsel(P::eq2(rv.z, mp), P::eq2(rv.w, mp), 0x0f000f00u), 0x00ff00ffu);
                    const uint32_t tl = t & 0xffffu, th = t >> 16;
                    const int el = tl ? ((__ffs(tl) - 1) >> 2) * 2 : 99;
                    const int eh = th ? ((__ffs(th) - 1) >> 2) * 2 + 1 : 99;
                    e0 = min(el, eh);
                } else {
                    float f[VE];
                    E::unpack(rv, f);
                    e0 = 0;
#pragma unroll
                    for (int e = VE - 1; e >= 0; --e)
                        if (f[e] >= M) e0 = e;
                }
                cand = (lane + 32 * pick) * VE + e0;
                }
            }
        }
        const int pred = __reduce_min_sync(0xffffffffu, cand);
        const int rank = __reduce_add_sync(0xffffffffu, cnt) / P::kScale;
        if (lane == 0) {
            c1 += (rank < 1) ? 1u : 0u;
            ck += (rank < k) ? 1u : 0u;
            if (SCM) atomicAdd(&s_cm[y * C + pred], 1u);
            else if (confusion) atomicAdd(&confusion[(int64_t)y * C + pred], 1ULL);
        }
    }
    if (lane == 0) {
        if (c1) atomicAdd(&s_cnt[0], c1);
        if (ck) atomicAdd(&s_cnt[1], ck);
        if (inv) atomicAdd(&s_cnt[2], inv);
    }
    __syncthreads();
    if (SCM) {
        for (int i = threadIdx.x; i < C * C; i += kMxThreads) {
            const unsigned int v = s_cm[i];
            if (v) atomicAdd(&confusion[i], (unsigned long long)v);
        }
    }
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) atomicAdd(&counters[0], (unsigned long long)N);
        if (s_cnt[0]) atomicAdd(&counters[1], (unsigned long long)s_cnt[0]);
        if (s_cnt[1]) atomicAdd(&counters[2], (unsigned long long)s_cnt[1]);
        if (s_cnt[2]) atomicAdd(&counters[3], (unsigned long long)s_cnt[2]);
    }
}

// Reductions over the 16 lanes of a half-warp (xor butterflies stay inside the half; `m` = the half's lanes).
__device__ __forceinline__ int tb_half_max(unsigned m, int v) {
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) v = max(v, __shfl_xor_sync(m, v, o));
    return v;
}
__device__ __forceinline__ int tb_half_min(unsigned m, int v) {
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) v = min(v, __shfl_xor_sync(m, v, o));
    return v;
}
__device__ __forceinline__ int tb_half_add(unsigned m, int v) {
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) v += __shfl_xor_sync(m, v, o);
    return v;
}

// 16-bit logits: TWO rows per warp, one per half-warp.  A 2 KB row is only four vectors per lane of a full warp, and the
// work that does not depend on the row's length (label, thresholds, three reductions, the argmax search, the counters)
// was three quarters of a row's 273 instructions; a half-warp per row does that work for two rows at once.
template <int DT, bool SCM>
__global__ void __launch_bounds__(kMxThreads, 4) tb_metrics_row2_kernel(const void* __restrict__ logits, int64_t N, int C,
                                                                    int64_t row_stride, const int64_t* __restrict__ labels,
                                                                    int k, unsigned long long* __restrict__ counters,
                                                                    unsigned long long* __restrict__ confusion) {
    using E = Elem<DT>;
    using P = Packed<DT>;
    constexpr int VE = E::VE;
    constexpr int UN = 64 / VE;      // vectors per lane of a half-warp: 8 for 16-bit types (1024 elements per row at most)
    constexpr uint32_t kFill = DT == TB_F32 ? 0xff800000u : (DT == TB_BF16 ? 0xff80ff80u : 0xfc00fc00u);     // -inf
    __shared__ unsigned int s_cnt[3];
    extern __shared__ unsigned int s_cm[];        // SCM only: C * C bins
    if (threadIdx.x < 3) s_cnt[threadIdx.x] = 0;
    if (SCM) {
        for (int i = threadIdx.x; i < C * C; i += kMxThreads) s_cm[i] = 0;
    }
    __syncthreads();
    tb_launch_dependents();
    tb_grid_dependency_wait();

    // `lane` is the lane within the HALF-warp, `warp0` / `nwarps` the first row and the row stride of this half-warp
    const int lane = threadIdx.x & 15;
    const unsigned hmask = 0xffffu << (threadIdx.x & 16);
    const int64_t warp0 = ((int64_t)blockIdx.x * kMxWarps + (threadIdx.x >> 5)) * 2 + ((threadIdx.x >> 4) & 1);
    const int64_t nwarps = (int64_t)gridDim.x * kMxWarps * 2;
    const int nv = C / VE;
    unsigned int c1 = 0, ck = 0, inv = 0;
    bool take[UN], whole[UN];       // this lane has vector i; every lane has vector i (no -inf fill needed)
#pragma unroll
    for (int i = 0; i < UN; ++i) { take[i] = lane + 16 * i < nv; whole[i] = 16 * (i + 1) <= nv; }

    long long y64 = warp0 < N ? labels[warp0] : -1;
    const typename E::type* row = (const typename E::type*)logits + warp0 * row_stride;
    const int64_t row_step = nwarps * row_stride;
    for (int64_t n = warp0; n < N; n += nwarps, row += row_step) {
        const long long ycur = y64;
        if (n + nwarps < N) y64 = labels[n + nwarps];
        if (ycur < 0 || ycur >= C) { inv += 1; continue; }
        const int y = (int)ycur;
        const uint4* vp = reinterpret_cast<const uint4*>(row) + lane;
        uint4 r[UN];
#pragma unroll
        for (int i = 0; i < UN; ++i) {
            if (whole[i]) r[i] = tb_ldcs_now(vp + 16 * i);
            else r[i] = tb_ldcs_or(vp + 16 * i, take[i], kFill);
        }
        const typename E::type raw_y = row[y];
        // the elements that share the label's 16-byte vector and lie before it, one per lane (ties among them count)
        const int d0 = y & (VE - 1);
        const typename E::type raw_t = row[y - d0 + min(lane, VE - 1)];
        const float ly = E::to_float(raw_y);
        if (ly != ly) {
            // rare: the label's own logit is NaN -- the key route of the general kernel, element by element
            RowAcc<DT> a;
            a.init(raw_y, y);
            for (int j = lane; j < C; j += 16) a.one(E::get(row, j), j);
            const int mx = tb_half_max(hmask, a.bestk);
            const int pred = tb_half_min(hmask, a.bestk == mx ? a.bj0 : INT_MAX);
            const int rank = tb_half_add(hmask, a.cnt) / P::kScale;
            if (lane == 0) {
                c1 += (rank < 1) ? 1u : 0u;
                ck += (rank < k) ? 1u : 0u;
                if (SCM) atomicAdd(&s_cm[y * C + pred], 1u);
                else if (confusion) atomicAdd(&confusion[(int64_t)y * C + pred], 1ULL);
            }
            continue;
        }

        // ---- thresholds: x > hi  <=>  x outranks the label from behind it; x > lo  <=>  x >= l_y (from before it).
        // A vector wholly before the label is counted against lo, every other one against hi; the ties before the label
        // inside its own vector are the (lane < d0) term.
        int cnt = (lane < d0 && E::to_float(raw_t) == ly) ? P::kScale : 0;
        float m[UN];
        const int dl = y - lane * VE;                       // label position relative to this lane's first vector
        if constexpr (P::kOn) {
            uint32_t c = (uint32_t)raw_y;
            if (c == 0x8000u) c = 0;                                   // -0 == +0
            uint32_t l;
            if (c == P::kNegInf) l = 0x7fffu;                          // >= -inf: everything counts (NaN threshold: unordered)
            else if (c == 0u) l = 0x8001u;                             // just below zero
            else l = (c & 0x8000u) ? c + 1u : c - 1u;
            const uint32_t hi2 = c * 0x10001u, lo2 = l * 0x10001u;
#pragma unroll
            for (int i = 0; i < UN; ++i) {
                const int d = dl - 16 * i * VE;                        // >= VE: the whole vector lies before the label
                const uint32_t thr = d >= VE ? lo2 : hi2;
                const uint32_t t = tb_bitsel(tb_bitsel(P::gtu2(r[i].x, thr), P::gtu2(r[i].y, thr), 0x000f000fu),
                                             tb_bitsel(P::gtu2(r[i].z, thr), P::gtu2(r[i].w, thr), 0x0f000f00u), 0x00ff00ffu);
                cnt += __popc(t);
                const uint32_t mm = P::max2(P::max2(r[i].x, r[i].y), P::max2(r[i].z, r[i].w));
                m[i] = tb_max_nan(P::lo(mm), P::hi(mm));
            }
        } else {
            const float cc = ly + 0.0f;
            const uint32_t b = __float_as_uint(cc);
            const float hi = cc;
            float lo;
            if (b == 0xff800000u) lo = __uint_as_float(0x7fffffffu);
            else if (b == 0u) lo = __uint_as_float(0x80000001u);
            else lo = __uint_as_float((b & 0x80000000u) ? b + 1u : b - 1u);
#pragma unroll
            for (int i = 0; i < UN; ++i) {
                const int d = dl - 16 * i * VE;
                const float thr = d >= VE ? lo : hi;
                const float f0 = __uint_as_float(r[i].x), f1 = __uint_as_float(r[i].y), f2 = __uint_as_float(r[i].z),
                            f3 = __uint_as_float(r[i].w);
                cnt = cnt - tb_gtu(f0, thr) - tb_gtu(f1, thr);
                cnt = cnt - tb_gtu(f2, thr) - tb_gtu(f3, thr);
                m[i] = tb_max_nan(tb_max3_nan(f0, f1, f2), f3);
            }
        }
        // ---- argmax: warp maximum of the keys, then the first index among the lanes that hold it
        float M = m[0];
#pragma unroll
        for (int i = 1; i < UN; ++i) M = tb_max_nan(M, m[i]);
        const int key = tb_key(M);
        const int mx = tb_half_max(hmask, key);
        int cand = INT_MAX;
        if (key == mx) {
            if (M != M) {
                // a NaN is the maximum (rare): first NaN of this lane
                int pick = 0;
#pragma unroll
                for (int i = UN - 1; i >= 0; --i)
                    if (m[i] != m[i]) pick = i;
                if (lane + 16 * pick < nv) {
                    const uint4 rv = __ldg(vp + 16 * pick);
                    float f[VE];
                    E::unpack(rv, f);
                    int e0 = 0;
#pragma unroll
                    for (int e = VE - 1; e >= 0; --e)
                        if (f[e] != f[e]) e0 = e;
                    cand = (lane + 16 * pick) * VE + e0;
                }
            } else {
                int pick = 0;
#pragma unroll
                for (int i = UN - 1; i >= 0; --i)
                    if (m[i] >= M) pick = i;
                // The winning vector is read again (one 16-byte load by one lane, an L2 hit) instead of being kept:
                // holding all UN vectors for this moment cost the fp32 kernel its registers -- the compiler reused load
                // registers and a row took three DRAM round trips instead of one.
                if (lane + 16 * pick < nv) {
                const uint4 rv = __ldg(vp + 16 * pick);
                int e0;
                if constexpr (P::kOn) {
                    // packed equality against the maximum; nibble w of the low / high half <- word w's low / high element
                    const uint32_t mp = P::bits(M) * 0x10001u;
                    const uint32_t t = tb_bitsel(tb_bitsel(P::eq2(rv.x, mp), P::eq2(rv.y, mp), 0x000f000fu),
                                                 tb_bitsel(P::eq2(rv.z, mp), P::eq2(rv.w, mp), 0x0f000f00u), 0x00ff00ffu);
                    const uint32_t tl = t & 0xffffu, th = t >> 16;
                    const int el = tl ? ((__ffs(tl) - 1) >> 2) * 2 : 99;
                    const int eh = th ? ((__ffs(th) - 1) >> 2) * 2 + 1 : 99;
                    e0 = min(el, eh);
                } else {
                    float f[VE];
                    E::unpack(rv, f);
                    e0 = 0;
#pragma unroll
                    for (int e = VE - 1; e >= 0; --e)
                        if (f[e] >= M) e0 = e;
                }
                cand = (lane + 16 * pick) * VE + e0;
                }
            }
        }
        const int pred = tb_half_min(hmask, cand);
        const int rank = tb_half_add(hmask, cnt) / P::kScale;
        if (lane == 0) {
            c1 += (rank < 1) ? 1u : 0u;
            ck += (rank < k) ? 1u : 0u;
            if (SCM) atomicAdd(&s_cm[y * C + pred], 1u);
            else if (confusion) atomicAdd(&confusion[(int64_t)y * C + pred], 1ULL);
        }
    }
    if (lane == 0) {
        if (c1) atomicAdd(&s_cnt[0], c1);
        if (ck) atomicAdd(&s_cnt[1], ck);
        if (inv) atomicAdd(&s_cnt[2], inv);
    }
    __syncthreads();
    if (SCM) {
        for (int i = threadIdx.x; i < C * C; i += kMxThreads) {
            const unsigned int v = s_cm[i];
            if (v) atomicAdd(&confusion[i], (unsigned long long)v);
        }
    }
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

// ---- int32 wire format of the packed metric state (the cross-rank exchange) -------------------------------
// state : int64 [4 + bins]  = n, correct_1, correct_k, invalid, cm...
// wire  : int32 [12 + bins] = the four header values as three 21-bit limbs each (sums over up to 1024 ranks stay below
//         2^31 and carry nothing from limb to limb, so the int64 totals are rebuilt exactly), then the bins as int32.
constexpr int kWireHeader = 12;
__global__ void __launch_bounds__(256) tb_state_pack_kernel(const long long* __restrict__ state, long long bins,
                                                            int* __restrict__ wire) {
    tb_grid_dependency_wait();
    const long long i0 = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i0 < 4) {
        const unsigned long long v = (unsigned long long)state[i0];
        wire[3 * i0] = (int)(v & 0x1fffffu);
        wire[3 * i0 + 1] = (int)((v >> 21) & 0x1fffffu);
        wire[3 * i0 + 2] = (int)((v >> 42) & 0x1fffffu);
    }
    for (long long i = i0; i < bins; i += (long long)gridDim.x * 256) wire[kWireHeader + i] = (int)state[4 + i];
}
__global__ void __launch_bounds__(256) tb_state_unpack_kernel(const int* __restrict__ wire, long long bins,
                                                              long long* __restrict__ state) {
    tb_grid_dependency_wait();
    const long long i0 = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i0 < 4)
        state[i0] = (long long)wire[3 * i0] + ((long long)wire[3 * i0 + 1] << 21) + ((long long)wire[3 * i0 + 2] << 42);
    for (long long i = i0; i < bins; i += (long long)gridDim.x * 256) state[4 + i] = (long long)wire[kWireHeader + i];
}

}  // namespace

extern "C" int tb_metric_state_pack_i32(const int64_t* state, int64_t bins, int32_t* wire, tb_stream_t stream) {
    TB_REQUIRE(bins >= 0, "tb_metric_state_pack_i32: negative bins");
    TB_REQUIRE(state && wire, "tb_metric_state_pack_i32: NULL argument");
    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;
    int64_t blocks = (bins + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > (int64_t)dev.sm_count * 8) blocks = (int64_t)dev.sm_count * 8;
    tb_state_pack_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const long long*)state, (long long)bins, wire);
    TB_LAUNCH_CHECK("tb_state_pack_kernel");
    return TB_OK;
}

extern "C" int tb_metric_state_unpack_i32(const int32_t* wire, int64_t bins, int64_t* state, tb_stream_t stream) {
    TB_REQUIRE(bins >= 0, "tb_metric_state_unpack_i32: negative bins");
    TB_REQUIRE(state && wire, "tb_metric_state_unpack_i32: NULL argument");
    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;
    int64_t blocks = (bins + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > (int64_t)dev.sm_count * 8) blocks = (int64_t)dev.sm_count * 8;
    tb_state_unpack_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(wire, (long long)bins, (long long*)state);
    TB_LAUNCH_CHECK("tb_state_unpack_kernel");
    return TB_OK;
}

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
    auto* cnt = reinterpret_cast<unsigned long long*>(counters);
    auto* cm = reinterpret_cast<unsigned long long*>(confusion);
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kMxThreads);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;

    // few classes: bins in shared memory (see the kernel); more than 24 KB of them leaves room for 4 CTAs per SM
    const size_t scm_bytes = (size_t)C * (size_t)C * 4;
    const bool scm = confusion != nullptr && scm_bytes <= (size_t)kScmMaxBytes;
    cfg.dynamicSmemBytes = scm ? scm_bytes : 0;
    // 16-bit rows of one batch: the two-rows-per-warp kernel (two resident CTAs per SM, 16 rows per CTA round)
    const int ve0 = 16 / es;
    const bool row2 = es == 2 && ((uintptr_t)logits % 16) == 0 && ((row_stride * es) % 16) == 0 && (C % ve0) == 0 && C / ve0 <= 128;
    if (row2) blocks = (N + 2 * kMxWarps - 1) / (2 * kMxWarps);
    const int64_t cap = (int64_t)dev.sm_count * (row2 ? 8 : ((scm && scm_bytes > 24 * 1024) ? 4 : 8));   // two rounds of the resident CTAs
    if (blocks > cap) blocks = cap;
    TB_REQUIRE(N / (blocks * kMxWarps) < ((int64_t)1 << 28), "tb_metrics_update: N too large for one launch");
    cfg.gridDim = dim3((unsigned)blocks);
    const bool aligned = ((uintptr_t)logits % 16) == 0 && ((row_stride * es) % 16) == 0;
#define TB_MX_LAUNCH2(DT, AL)                                                                                          \
    do {                                                                                                               \
        if (scm) {                                                                                                     \
            TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_kernel<DT, AL, true>, logits, N, C, row_stride, labels, k, cnt, cm));  \
        } else {                                                                                                       \
            TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_kernel<DT, AL, false>, logits, N, C, row_stride, labels, k, cnt, cm)); \
        }                                                                                                              \
    } while (0)
#define TB_MX_LAUNCH(DT)                \
    do {                                \
        if (aligned) TB_MX_LAUNCH2(DT, true);  \
        else TB_MX_LAUNCH2(DT, false);         \
    } while (0)
    // rows of one batch (aligned, whole vectors, at most 1024 classes): the lean kernel; everything else (odd widths,
    // misaligned or longer rows) the general one
    const int ve = 16 / es;
    const bool one_batch = aligned && (C % ve) == 0 && C / ve <= 32 * (32 / ve);
#define TB_MX_ROW(DT)                                                                                                         \
    do {                                                                                                                      \
        if (scm) {                                                                                                            \
            TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_row_kernel<DT, true>, logits, N, C, row_stride, labels, k, cnt, cm));     \
        } else {                                                                                                              \
            TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_row_kernel<DT, false>, logits, N, C, row_stride, labels, k, cnt, cm));    \
        }                                                                                                                     \
    } while (0)
#define TB_MX_ROW2(DT)                                                                                                        \
    do {                                                                                                                      \
        if (scm) {                                                                                                            \
            TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_row2_kernel<DT, true>, logits, N, C, row_stride, labels, k, cnt, cm));    \
        } else {                                                                                                              \
            TB_CUDA(cudaLaunchKernelEx(&cfg, tb_metrics_row2_kernel<DT, false>, logits, N, C, row_stride, labels, k, cnt, cm));   \
        }                                                                                                                     \
    } while (0)
    if (one_batch) {
        if (logits_dtype == TB_F32) TB_MX_ROW(TB_F32);
        else if (logits_dtype == TB_BF16) TB_MX_ROW2(TB_BF16);
        else TB_MX_ROW2(TB_F16);
        TB_LAUNCH_CHECK("tb_metrics_row_kernel");
        return TB_OK;
    }
#undef TB_MX_ROW2
#undef TB_MX_ROW
    if (logits_dtype == TB_F32) TB_MX_LAUNCH(TB_F32);
    else if (logits_dtype == TB_BF16) TB_MX_LAUNCH(TB_BF16);
    else TB_MX_LAUNCH(TB_F16);
#undef TB_MX_LAUNCH
#undef TB_MX_LAUNCH2
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
