// K3 + K4a — limb scoring of every candidate pair, ordering, greedy connection.  One CTA per (image, limb).
//
// Reference: utils/pafprocess/pafprocess.cpp:52-130 with get_paf_scores (:311-327), round2int (:329-331) and
// comp_candidate (:333-335).  The arithmetic below reproduces the reference's mix of int, float and double step
// by step; every float operation is an explicit round-to-nearest intrinsic so that no
// FMA contraction can creep in (the reference build has none).
//
// Limb values come from one of two sources:
//   IBP_LIMB_FULLRES_HWC  the caller's (H,W,C) map, nearest-pixel lookup exactly like the PAF macro (:9, :322);
//   IBP_LIMB_FEAT_PLANAR  the feature-resolution plane of this limb staged in shared memory by the bulk-copy engine
//                         (h*w*4 bytes, one contiguous run); each sample evaluates OpenCV's bicubic x`up` at that
//                         pixel, which equals indexing cv2.resize(pafs, fx=up, fy=up, INTER_CUBIC) (evaluate.py:77-80)
//                         bit for bit without ever writing the 31 MB up-sampled map.
//
// Ordering: candidates are compacted in (a-major, b-minor) pair order, then sorted by a bitonic network on
// (score descending, candidate index descending).  That equals libstdc++'s std::sort with the reference's non-strict
// `>=` comparator for n <= 16 (insertion sort: ties end up in reverse input order) and for any n without duplicate
// keys.  A limb with n > 16 AND duplicate keys is re-ordered by one thread running the introsort emulation of
// stl_sort_emul.h (status bit IBP_ST_TIE_EMULATED; IBP_ST_TIE_UNDEFINED when the real sort would have left the array).
#pragma once
#include "ibp_device.cuh"
#include "stl_sort_emul.h"

namespace ibp {

struct LimbArgs {
    const DevPeak *peaks;   // (batch, 18, cap)
    const int *part_count;  // (batch, 18)
    int cap;
    const float *src;       // limb values, see mode
    int mode;
    size_t image_stride;    // floats between images in src
    int H, W, C;            // mode 0: full-resolution dims and channel count (stride); sample clamp bounds
    int h, w;               // mode 1: feature dims
    int up;
    int stage_plane;        // mode 1: plane fits the shared buffer
    int min_img_size;
    float thr, step_ratio, w_limb, w_a, w_b;
    int max_steps;
    int cand_cap;           // power of two
    DevLink *links;         // (batch, 30, cap)
    int *link_count;        // (batch, 30)
    uint32_t *status;       // (batch)
};

constexpr int kLimbThreads = 256;

__device__ __forceinline__ uint32_t float_orderable(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ uint32_t float_orderable_inv(uint32_t k) { return (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; }

struct Sampler {
    const float *base;  // mode 0: image base; mode 1: plane base (shared or global)
    int mode, limb;
    int H, W, C, h, w, up;
    __device__ __forceinline__ float at(int px, int py) const {
        if (mode == IBP_LIMB_FULLRES_HWC) {
            px = clampi(px, 0, W - 1);  // the reference does not check (:322); stay inside the caller's buffer
            py = clampi(py, 0, H - 1);
            return __ldg(base + (size_t)C * ((size_t)px + (size_t)W * py) + limb);
        }
        px = clampi(px, 0, w * up - 1);
        py = clampi(py, 0, h * up - 1);
        return cubic_sample(base, h, w, up, px, py);
    }
};

__device__ __forceinline__ float limb_length(const DevPeak &A, const DevPeak &B) {
    const float dx = __int2float_rn(B.x - A.x), dy = __int2float_rn(B.y - A.y);  // :68-69
    return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));        // :70
}

// pafprocess.cpp:66-104 for one pair.  Returns true when both criteria hold.
__device__ __forceinline__ bool score_pair(const DevPeak &A, const DevPeak &B, const Sampler &s, const LimbArgs &a, float &limb_score,
                                           float &rank) {
    const float len = limb_length(A, B);
    if ((double)len < 1e-12) return false;                                                     // :71
    const int steps = min(__double2int_rz(__dadd_rn((double)__fadd_rn(len, 1.0f), 0.5)), a.max_steps);  // :73
    const float denom = __int2float_rn(steps - 1);
    const float sx = __fdiv_rn(__int2float_rn(B.x - A.x), denom);  // :315
    const float sy = __fdiv_rn(__int2float_rn(B.y - A.y), denom);  // :316
    const float ax = __int2float_rn(A.x), ay = __int2float_rn(A.y);
    float sum = 0.0f;
    int above = 0;
    for (int i = 0; i < steps; i++) {
        const float fi = __int2float_rn(i);
        const int px = __double2int_rz(__dadd_rn((double)__fadd_rn(ax, __fmul_rn(fi, sx)), 0.5));  // :319, :329-331
        const int py = __double2int_rz(__dadd_rn((double)__fadd_rn(ay, __fmul_rn(fi, sy)), 0.5));  // :320
        const float v = s.at(px, py);
        sum = __fadd_rn(sum, v);  // :85
        above += (v > a.thr);     // :87
    }
    // :90  float + min(0.0, 0.5*min_img_size/len - 1.0) evaluated in double, narrowed on assignment
    const double pen_raw = __dsub_rn(__ddiv_rn(__dmul_rn(0.5, (double)a.min_img_size), (double)len), 1.0);
    const double pen = (pen_raw < 0.0) ? pen_raw : 0.0;
    limb_score = __double2float_rn(__dadd_rn((double)__fdiv_rn(sum, __int2float_rn(steps)), pen));
    const float need = __fmul_rn(__int2float_rn(steps), a.step_ratio);  // :91
    if (!(__int2float_rn(above) > need && limb_score > 0.0f)) return false;  // :93
    rank = __fadd_rn(__fadd_rn(__fmul_rn(a.w_limb, limb_score), __fmul_rn(a.w_a, A.score)), __fmul_rn(a.w_b, B.score));  // :94-96
    return true;
}

__global__ void __launch_bounds__(kLimbThreads) limbs_kernel(LimbArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int limb = blockIdx.x, img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int pa = c_limb_a[limb], pb = c_limb_b[limb];
    const int nA = a.part_count[img * kParts + pa], nB = a.part_count[img * kParts + pb];
    int *link_count = a.link_count + img * kLimbs + limb;
    if (nA == 0 || nB == 0) {  // no pair => no candidate => no connection (covers the both-empty skip at :57)
        if (tid == 0) *link_count = 0;
        return;
    }

    // ---- carve shared memory ----
    // region 0: limb plane while scoring (mode 1, staged), re-used for the introsort emulation array afterwards
    size_t off = 0;
    float *plane_s = reinterpret_cast<float *>(smem_raw);
    {
        size_t plane_bytes = (a.mode == IBP_LIMB_FEAT_PLANAR && a.stage_plane) ? (size_t)a.h * a.w * 4 : 0;
        size_t emu_bytes = (size_t)a.cand_cap * 8;
        off = plane_bytes > emu_bytes ? plane_bytes : emu_bytes;
        off = (off + 15) & ~(size_t)15;
    }
    uint64_t *emu = reinterpret_cast<uint64_t *>(smem_raw);
    uint64_t *keys = reinterpret_cast<uint64_t *>(smem_raw + off);
    off += (size_t)a.cand_cap * 8;
    float *cand_limb = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)a.cand_cap * 4;
    uint32_t *cand_pair = reinterpret_cast<uint32_t *>(smem_raw + off);
    off += (size_t)a.cand_cap * 4;
    DevPeak *sA = reinterpret_cast<DevPeak *>(smem_raw + off);
    off += (size_t)a.cap * sizeof(DevPeak);
    DevPeak *sB = reinterpret_cast<DevPeak *>(smem_raw + off);
    off += (size_t)a.cap * sizeof(DevPeak);
    unsigned char *used_a = smem_raw + off;
    off += (size_t)a.cap;
    unsigned char *used_b = smem_raw + off;
    __shared__ uint64_t bar;
    __shared__ int wcnt[kLimbThreads / 32];
    __shared__ int s_flag;

    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
        s_flag = 0;
    }
    const DevPeak *gA = a.peaks + ((size_t)img * kParts + pa) * a.cap;
    const DevPeak *gB = a.peaks + ((size_t)img * kParts + pb) * a.cap;
    for (int i = tid; i < nA; i += kLimbThreads) {
        sA[i] = gA[i];
        used_a[i] = 0;
    }
    for (int i = tid; i < nB; i += kLimbThreads) {
        sB[i] = gB[i];
        used_b[i] = 0;
    }
    __syncthreads();

    Sampler smp;
    smp.mode = a.mode;
    smp.limb = limb;
    smp.H = a.H;
    smp.W = a.W;
    smp.C = a.C;
    smp.h = a.h;
    smp.w = a.w;
    smp.up = a.up;
    if (a.mode == IBP_LIMB_FEAT_PLANAR) {
        const float *gplane = a.src + (size_t)img * a.image_stride + (size_t)limb * a.h * a.w;
        if (a.stage_plane) {
            uint32_t phase = 0;
            rows_to_smem(plane_s, gplane, (uint32_t)(a.h * a.w), &bar, phase);
            smp.base = plane_s;
        } else {
            smp.base = gplane;
        }
    } else {
        smp.base = a.src + (size_t)img * a.image_stride;
    }

    // ---- score all pairs; compact the accepted ones in pair order ----
    const int npairs = nA * nB;
    int ncand = 0;
    for (int base = 0; base < npairs; base += kLimbThreads) {
        const int p = base + tid;
        bool ok = false;
        float ls = 0.f, rk = 0.f;
        int ia = 0, ib = 0;
        if (p < npairs) {
            ia = p / nB;
            ib = p - ia * nB;
            ok = score_pair(sA[ia], sB[ib], smp, a, ls, rk);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) wcnt[warp] = __popc(bal);
        __syncthreads();
        int woff = 0, tot = 0;
#pragma unroll
        for (int i = 0; i < kLimbThreads / 32; i++) {
            const int t = wcnt[i];
            if (i < warp) woff += t;
            tot += t;
        }
        if (ok) {
            const int idx = ncand + woff + __popc(bal & ((1u << lane) - 1u));
            if (idx < a.cand_cap) {
                keys[idx] = ((uint64_t)float_orderable(rk) << 32) | (uint32_t)idx;
                cand_limb[idx] = ls;
                cand_pair[idx] = ((uint32_t)ia << 16) | (uint32_t)ib;
            }
        }
        ncand += tot;
        __syncthreads();
    }
    if (ncand > a.cand_cap) {
        if (tid == 0) atomicOr(a.status + img, IBP_ST_CAND_OVERFLOW);
        ncand = a.cand_cap;
    }
    if (ncand == 0) {
        if (tid == 0) *link_count = 0;
        return;
    }

    // ---- order: bitonic network, descending on (score, index) ----
    int np2 = 1;
    while (np2 < ncand) np2 <<= 1;
    for (int i = ncand + tid; i < np2; i += kLimbThreads) keys[i] = 0;  // below every real key
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += kLimbThreads) {
                const int l = i ^ j;
                if (l > i) {
                    const uint64_t x = keys[i], y = keys[l];
                    const bool desc = ((i & k) == 0);
                    if (desc ? (x < y) : (x > y)) {
                        keys[i] = y;
                        keys[l] = x;
                    }
                }
            }
            __syncthreads();
        }
    }
    // duplicate keys?
    bool dup = false;
    for (int i = tid; i + 1 < ncand; i += kLimbThreads) dup |= ((uint32_t)(keys[i] >> 32) == (uint32_t)(keys[i + 1] >> 32));
    const bool emulate = __syncthreads_or(dup && ncand > 16);
    if (emulate) {
        // rebuild the array in original (pair) order with raw float bits, then walk libstdc++'s steps on one thread
        for (int i = tid; i < ncand; i += kLimbThreads) {
            const uint64_t kv = keys[i];
            const uint32_t idx = (uint32_t)kv;
            emu[idx] = ((uint64_t)float_orderable_inv((uint32_t)(kv >> 32)) << 32) | idx;
        }
        __syncthreads();
        if (tid == 0) {
            const bool defined = ibp_sortemu::sort_like_libstdcxx(emu, ncand);
            atomicOr(a.status + img, defined ? IBP_ST_TIE_EMULATED : (IBP_ST_TIE_EMULATED | IBP_ST_TIE_UNDEFINED));
        }
        __syncthreads();
    }
    const uint64_t *order = emulate ? emu : keys;

    // ---- greedy connection (:111-129), warp 0: 32 candidates per step, leaders taken in order ----
    if (warp == 0) {
        const int max_links = min(nA, nB);
        DevLink *out = a.links + ((size_t)img * kLimbs + limb) * a.cap;
        int nl = 0;
        for (int base = 0; base < ncand && nl < max_links; base += 32) {
            const int i = base + lane;
            bool valid = i < ncand;
            int ia = 0, ib = 0, cidx = 0;
            if (valid) {
                cidx = (int)(uint32_t)order[i];
                const uint32_t pr = cand_pair[cidx];
                ia = (int)(pr >> 16);
                ib = (int)(pr & 0xffffu);
            }
            while (nl < max_links) {
                const bool free_now = valid && !used_a[ia] && !used_b[ib];
                const uint32_t bal = __ballot_sync(0xffffffffu, free_now);
                if (bal == 0) break;
                const int leader = __ffs(bal) - 1;
                if (lane == leader) {
                    used_a[ia] = 1;
                    used_b[ib] = 1;
                    DevLink l;
                    l.id_a = sA[ia].id;
                    l.id_b = sB[ib].id;
                    l.limb = cand_limb[cidx];
                    l.len = limb_length(sA[ia], sB[ib]);
                    out[nl] = l;
                    valid = false;
                }
                nl++;
                __syncwarp();
            }
        }
        if (lane == 0) *link_count = nl;
    }
}

}  // namespace ibp
