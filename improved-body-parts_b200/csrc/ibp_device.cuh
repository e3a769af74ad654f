// Device-side building blocks shared by the kernels: constant tables, exact-order float helpers, the OpenCV-order
// bicubic taps, and the 1-D bulk (TMA) global->shared copy with its mbarrier.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/ibp.h"

namespace ibp {

constexpr int kParts = IBP_NUM_PARTS;
constexpr int kLimbs = IBP_NUM_LIMBS;
constexpr int kLayers = IBP_NUM_LAYERS;

struct DevPeak {  // 16 B — reference Peak, pafprocess.h:29-34
    int x, y;
    float score;
    int id;
};
struct __align__(16) DevLink {  // 32 B — the fields of Connection (pafprocess.h:60-67) that grouping reads, plus ...
    int id_a, id_b;
    float limb;
    float len;
    float score_a, score_b;  // ... the two peak scores grouping would fetch through peak_infos_line (pafprocess.cpp:160, :261-262)
    float gate;              // ... and (float)(int)(len * LIMB_LENGTH_RATE), the connection's share of the :154 / :186 test
    float pad;
};

// limb end points: pafprocess.h:20-27 COCOPAIRS == config/config.py:114-121
__constant__ int c_limb_a[kLimbs] = {1, 1, 1, 1, 1, 0, 0, 14, 15, 1, 2, 3, 1, 5, 6, 1, 8, 9, 1, 11, 12, 0, 0, 2, 8, 5, 11, 16, 17, 8};
__constant__ int c_limb_b[kLimbs] = {0, 14, 15, 16, 17, 14, 15, 16, 17, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 2, 5, 8, 12, 11, 9, 2, 5, 11};
// channel read from the flipped image for output channel c: config/config.py:150-152 (flip_paf_ord, then
// 30 + flip_heat_ord)
__constant__ int c_flip_src[kLayers] = {0, 2, 1, 4, 3, 6, 5, 8, 7, 12, 13, 14, 9, 10, 11, 18, 19, 20, 15, 16, 17, 22, 21, 25, 26, 23, 24, 28, 27, 29,
                                        30, 31, 35, 36, 37, 32, 33, 34, 41, 42, 43, 38, 39, 40, 45, 44, 47, 46, 48, 49};

// ---------------------------------------------------------------------------------------------------------------
// OpenCV cubic taps (cv2.resize INTER_CUBIC, A = -0.75).  The library is compiled with -fmad=false, and the
// intrinsics below pin the evaluation order on top of that: the reference's numbers come from code without FMA
// contraction (OpenCV baseline SSE3; pafprocess built by distutils without -march).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cubic_weights(float t, float *w) {
    const float A = -0.75f;
    const float t1 = __fadd_rn(t, 1.0f);
    const float u = __fsub_rn(1.0f, t);
    // coeffs[0] = ((A*(x+1) - 5*A)*(x+1) + 8*A)*(x+1) - 4*A
    w[0] = __fsub_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fsub_rn(__fmul_rn(A, t1), 5.0f * A), t1), 8.0f * A), t1), 4.0f * A);
    // coeffs[1] = ((A+2)*x - (A+3))*x*x + 1
    w[1] = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(A + 2.0f, t), A + 3.0f), t), t), 1.0f);
    // coeffs[2] = ((A+2)*(1-x) - (A+3))*(1-x)*(1-x) + 1
    w[2] = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(A + 2.0f, u), A + 3.0f), u), u), 1.0f);
    // coeffs[3] = 1 - coeffs[0] - coeffs[1] - coeffs[2]
    w[3] = __fsub_rn(__fsub_rn(__fsub_rn(1.0f, w[0]), w[1]), w[2]);
}

// destination index d of an n_src -> n_src*up axis: first source index (unclamped) and the four weights
__device__ __forceinline__ int cubic_tap(int d, int up, float *w) {
    if (up == 4) {
        // (d+0.5)/4 - 0.5 is exact: floor = (d-2)>>2, fraction = ((d-2)&3)/4 + 1/8
        const int s = (d - 2) >> 2;
        const float t = __fadd_rn(__int2float_rn((d - 2) & 3) * 0.25f, 0.125f);
        cubic_weights(t, w);
        return s;
    }
    const double scale = __ddiv_rn(1.0, (double)up);
    float f = __double2float_rn(__dsub_rn(__dmul_rn(__dadd_rn((double)d, 0.5), scale), 0.5));
    const int s = __float2int_rd(f);
    f = __fsub_rn(f, __int2float_rn(s));
    cubic_weights(f, w);
    return s;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// For up == 4 the four weights depend only on the phase (d-2)&3 (fractions 1/8, 3/8, 5/8, 7/8): a 4x4 table in
// shared memory, filled with the very same cubic_weights() evaluation so that table and on-the-fly values are the
// same bits.  Call from all threads; the caller synchronises before the first use.
__device__ __forceinline__ void cubic_tab4_init(float (*tab)[4]) {
    if (threadIdx.x < 4) cubic_weights(__fadd_rn(__int2float_rn(threadIdx.x) * 0.25f, 0.125f), tab[threadIdx.x]);
}

// horizontal 4-tap: ((s0*a0 + s1*a1) + s2*a2) + s3*a3   (OpenCV HResizeCubic)
__device__ __forceinline__ float hsum4(float s0, float s1, float s2, float s3, const float *a) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(s0, a[0]), __fmul_rn(s1, a[1])), __fmul_rn(s2, a[2])), __fmul_rn(s3, a[3]));
}
// vertical 4-tap: s0*b0 + (s1*b1 + (s2*b2 + s3*b3))   (OpenCV VResizeCubicVec_32f)
__device__ __forceinline__ float vsum4(float s0, float s1, float s2, float s3, const float *b) {
    return __fadd_rn(__fmul_rn(s0, b[0]), __fadd_rn(__fmul_rn(s1, b[1]), __fadd_rn(__fmul_rn(s2, b[2]), __fmul_rn(s3, b[3]))));
}

__device__ __forceinline__ float to_f32(float a) { return a; }
__device__ __forceinline__ float to_f32(__half a) { return __half2float(a); }

// value of pixel (X,Y) of cv2.resize(plane(h,w), fx=up, fy=up, INTER_CUBIC); `plane` may be shared or global.  A plane of
// __half holds values that are exact in half precision (a float16 flip average): widening them is lossless.
template <typename T>
__device__ __forceinline__ float cubic_sample(const T *plane, int h, int w, int up, int X, int Y) {
    float xw[4], yw[4];
    const int sx = cubic_tap(X, up, xw);
    const int sy = cubic_tap(Y, up, yw);
    const int x0 = clampi(sx - 1, 0, w - 1), x1 = clampi(sx, 0, w - 1), x2 = clampi(sx + 1, 0, w - 1), x3 = clampi(sx + 2, 0, w - 1);
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const T *row = plane + (size_t)clampi(sy - 1 + k, 0, h - 1) * w;
        r[k] = hsum4(to_f32(row[x0]), to_f32(row[x1]), to_f32(row[x2]), to_f32(row[x3]), xw);
    }
    return vsum4(r[0], r[1], r[2], r[3], yw);
}

// same value through the phase table (up == 4 only)
template <typename T>
__device__ __forceinline__ float cubic_sample4(const T *plane, int h, int w, const float (*tab)[4], int X, int Y) {
    const int sx = (X - 2) >> 2, sy = (Y - 2) >> 2;
    const float *xw = tab[(X - 2) & 3], *yw = tab[(Y - 2) & 3];
    const int x0 = clampi(sx - 1, 0, w - 1), x1 = clampi(sx, 0, w - 1), x2 = clampi(sx + 1, 0, w - 1), x3 = clampi(sx + 2, 0, w - 1);
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const T *row = plane + (size_t)clampi(sy - 1 + k, 0, h - 1) * w;
        r[k] = hsum4(to_f32(row[x0]), to_f32(row[x1]), to_f32(row[x2]), to_f32(row[x3]), xw);
    }
    return vsum4(r[0], r[1], r[2], r[3], yw);
}

// ---------------------------------------------------------------------------------------------------------------
// mbarrier + 1-D bulk copy (TMA engine, SASS UBLKCP): one elected thread moves a contiguous run of rows of a
// planar map into shared memory; everybody waits on the barrier's phase.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Programmatic dependent launch: let the next kernel of the stream be set up now / wait for the previous one's results.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void cp_async4(void *dst_smem, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}

// Cooperative "rows to shared": the bulk engine when source, destination and size are 16-byte aligned (every map
// the reference pads to a multiple of 64 pixels is), plain coalesced loads otherwise.  Must be called by all
// threads of the CTA; returns after the data is visible to all of them.  `phase` is the barrier's running parity.
__device__ __forceinline__ void rows_to_smem(float *dst, const float *src, uint32_t n_floats, uint64_t *bar, uint32_t &phase) {
    const uint32_t bytes = n_floats * 4u;
    const bool bulk_ok = ((bytes & 15u) == 0) && ((((uintptr_t)src) & 15u) == 0) && ((smem_u32(dst) & 15u) == 0);
    if (bulk_ok) {
        if (threadIdx.x == 0) {
            // chunks keep each transaction count well inside the mbarrier's tx-count range
            uint32_t done = 0;
            mbar_expect_tx(bar, bytes);
            while (done < bytes) {
                uint32_t chunk = bytes - done;
                if (chunk > 65536u) chunk = 65536u;
                bulk_g2s((char *)dst + done, (const char *)src + done, chunk, bar);
                done += chunk;
            }
        }
        mbar_wait(bar, phase);
        phase ^= 1u;
    } else {
        for (uint32_t i = threadIdx.x; i < n_floats; i += blockDim.x) dst[i] = __ldg(src + i);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Flip / channel-permutation average fused into the staging load (utils/parse_skeletons.py:91-93):
//     dst[r][x] = (up[row0+r][x] + twin[row0+r][w-1-x]) / 2      averaged in the element type T like numpy does.
// 128-bit loads of the upright run and of the mirrored run (reversed in registers), four independent pairs in
// flight per thread, 128-bit shared stores.  Falls back to scalar loads for widths / alignments that do not
// vectorise.  `twin == nullptr` copies (through the bulk-copy engine when T is float and everything is aligned).
// Must be called by all threads; returns after the rows are visible to the whole CTA.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float avg_like_numpy(float a, float b) { return __fmul_rn(__fadd_rn(a, b), 0.5f); }
__device__ __forceinline__ float avg_like_numpy(__half a, __half b) {
    // numpy float16: each ufunc rounds to half — (a+b) -> half, then /2 -> half
    const __half s = __float2half_rn(__fadd_rn(__half2float(a), __half2float(b)));
    return __half2float(__float2half_rn(__fmul_rn(__half2float(s), 0.5f)));
}
// one 16-byte vector of a staged row folded with the vector that mirrors it in the twin row: o[q] = avg(o[q], t[V-1-q]),
// kept in the element type (the float16 average is exact in float16)
__device__ __forceinline__ void fold_mirrored(float4 &o, const float4 &t) {
    o.x = avg_like_numpy(o.x, t.w);
    o.y = avg_like_numpy(o.y, t.z);
    o.z = avg_like_numpy(o.z, t.y);
    o.w = avg_like_numpy(o.w, t.x);
}
__device__ __forceinline__ void fold_mirrored(uint4 &o, const uint4 &t) {  // eight __half
    __half *a = reinterpret_cast<__half *>(&o);
    const __half *m = reinterpret_cast<const __half *>(&t);
#pragma unroll
    for (int q = 0; q < 8; q++) a[q] = __float2half_rn(avg_like_numpy(a[q], m[7 - q]));
}
template <typename T>
struct Vec16 {
    using type = float4;
};
template <>
struct Vec16<__half> {
    using type = uint4;
};

template <typename T>
__device__ __forceinline__ void stage_rows_fused(float *dst, const T *up, const T *twin, int row0, int nrows, int w, uint64_t *bar,
                                                 uint32_t &phase) {
    constexpr int V = 16 / (int)sizeof(T);  // elements per 128-bit load
    const T *up0 = up + (size_t)row0 * w;
    if (twin == nullptr && sizeof(T) == 4) {
        rows_to_smem(dst, reinterpret_cast<const float *>(up0), (uint32_t)(nrows * w), bar, phase);
        return;
    }
    const bool vec_ok = (w % V == 0) && ((((uintptr_t)up) & 15u) == 0) && (twin == nullptr || (((uintptr_t)twin) & 15u) == 0) &&
                        ((smem_u32(dst) & 15u) == 0);
    if (vec_ok) {
        const int wv = w / V, nvec = nrows * wv, T_ = blockDim.x;
        const uint4 *u4 = reinterpret_cast<const uint4 *>(up0);
        const uint4 *t4 = twin ? reinterpret_cast<const uint4 *>(twin + (size_t)row0 * w) : nullptr;
        float4 *d4 = reinterpret_cast<float4 *>(dst);
        constexpr int U = 4;
        for (int base = threadIdx.x; base < nvec; base += U * T_) {
            uint4 a[U], m[U];
#pragma unroll
            for (int k = 0; k < U; k++) {
                const int i = base + k * T_;
                if (i < nvec) {
                    a[k] = __ldcs(u4 + i);
                    if (t4) {
                        const int r = i / wv, c = i - r * wv;
                        m[k] = __ldcs(t4 + r * wv + (wv - 1 - c));
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < U; k++) {
                const int i = base + k * T_;
                if (i < nvec) {
                    const T *av = reinterpret_cast<const T *>(&a[k]);
                    const T *mv = reinterpret_cast<const T *>(&m[k]);
                    float o[V];
#pragma unroll
                    for (int q = 0; q < V; q++) o[q] = t4 ? avg_like_numpy(av[q], mv[V - 1 - q]) : to_f32(av[q]);
#pragma unroll
                    for (int q = 0; q < V / 4; q++) d4[(size_t)i * (V / 4) + q] = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
                }
            }
        }
    } else {
        const int n = nrows * w;
        const T *tw0 = twin ? twin + (size_t)row0 * w : nullptr;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            if (tw0) {
                const int r = i / w, c = i - r * w;
                dst[i] = avg_like_numpy(up0[i], tw0[r * w + (w - 1 - c)]);
            } else {
                dst[i] = to_f32(up0[i]);
            }
        }
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------------
// The same flip average as a bulk-copy (TMA) pipeline, float32 sources only.  The upright rows go straight to their
// final place in shared memory; the twin rows stream through a ring of NS slots of `slot_rows` rows each and are
// folded into the upright rows (mirrored) as they land, so up to nrows + NS*slot_rows rows are in flight per CTA
// without holding a single register.  bars[0] tracks the upright copy, bars[1..NS] the ring slots.
// flip_tma_ok() says whether the shapes/alignments allow it; flip_tma_issue() is called by ONE thread as early as
// possible; flip_tma_finish() by all threads, and returns once dst holds the fused rows for everybody.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool flip_tma_ok(const float *dst, const float *up, const float *twin, const float *ring, int w, int slot_rows) {
    return twin != nullptr && (w & 3) == 0 && slot_rows >= 1 && ((((uintptr_t)up) | ((uintptr_t)twin)) & 15u) == 0 &&
           ((smem_u32(dst) | smem_u32(ring)) & 15u) == 0;
}
// copies only; the caller has announced the bytes on `bar` (mbar_expect_tx)
__device__ __forceinline__ void bulk_chunks(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    uint32_t done = 0;
    while (done < bytes) {
        uint32_t chunk = bytes - done;
        if (chunk > 65536u) chunk = 65536u;
        bulk_g2s((char *)dst + done, (const char *)src + done, chunk, bar);
        done += chunk;
    }
}
__device__ __forceinline__ void bulk_rows(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    mbar_expect_tx(bar, bytes);
    uint32_t done = 0;
    while (done < bytes) {
        uint32_t chunk = bytes - done;
        if (chunk > 65536u) chunk = 65536u;
        bulk_g2s((char *)dst + done, (const char *)src + done, chunk, bar);
        done += chunk;
    }
}
template <int NS>
__device__ __forceinline__ void flip_tma_issue(float *dst, const float *up, const float *twin, int row0, int nrows, int w, float *ring,
                                               int slot_rows, uint64_t *bars) {
    bulk_rows(dst, up + (size_t)row0 * w, (uint32_t)nrows * w * 4u, &bars[0]);
    for (int s = 0; s < NS; s++) {
        const int first = s * slot_rows;
        if (first < nrows) {
            const int rows = min(slot_rows, nrows - first);
            bulk_rows(ring + (size_t)s * slot_rows * w, twin + (size_t)(row0 + first) * w, (uint32_t)rows * w * 4u, &bars[1 + s]);
        }
    }
}
template <int NS>
__device__ __forceinline__ void flip_tma_finish(float *dst, const float *twin, int row0, int nrows, int w, float *ring, int slot_rows,
                                                uint64_t *bars) {
    const int wv = w >> 2;
    const int shift = ((wv & (wv - 1)) == 0) ? (__ffs(wv) - 1) : -1;
    const int nchunks = (nrows + slot_rows - 1) / slot_rows;
    mbar_wait(&bars[0], 0);
    float4 *d4 = reinterpret_cast<float4 *>(dst);
    for (int c = 0; c < nchunks; c++) {
        const int s = c % NS;
        mbar_wait(&bars[1 + s], (uint32_t)((c / NS) & 1));
        const int first = c * slot_rows;
        const int rows = min(slot_rows, nrows - first);
        const float4 *r4 = reinterpret_cast<const float4 *>(ring + (size_t)s * slot_rows * w);
        float4 *o4 = d4 + (size_t)first * wv;
        for (int i = threadIdx.x; i < rows * wv; i += blockDim.x) {
            const int r = (shift >= 0) ? (i >> shift) : (i / wv);
            const int col = i - r * wv;
            float4 a = o4[i];
            const float4 m = r4[r * wv + (wv - 1 - col)];
            a.x = avg_like_numpy(a.x, m.w);
            a.y = avg_like_numpy(a.y, m.z);
            a.z = avg_like_numpy(a.z, m.y);
            a.w = avg_like_numpy(a.w, m.x);
            o4[i] = a;
        }
        if (c + NS < nchunks) {
            __syncthreads();  // everybody has read slot s
            if (threadIdx.x == 0) {
                fence_proxy_async();  // order those generic-proxy reads before the async-proxy refill
                const int nfirst = (c + NS) * slot_rows;
                const int nrows2 = min(slot_rows, nrows - nfirst);
                bulk_rows(ring + (size_t)s * slot_rows * w, twin + (size_t)(row0 + nfirst) * w, (uint32_t)nrows2 * w * 4u, &bars[1 + s]);
            }
        }
    }
    __syncthreads();
}

}  // namespace ibp
