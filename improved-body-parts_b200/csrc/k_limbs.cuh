// K3 + K4 — limb scoring of every candidate pair, ordering, greedy connection (one CTA per image and limb), then
// person grouping by the last CTA of each image.
//
// Reference: utils/pafprocess/pafprocess.cpp:52-130 with get_paf_scores (:311-327), round2int (:329-331) and
// comp_candidate (:333-335); grouping :133-282 (k_group.cuh).  The arithmetic below reproduces the reference's mix
// of int, float and double step by step; every float operation is an explicit
// round-to-nearest intrinsic so that no FMA contraction can creep in (the reference build has none).
//
// Limb values come from one of three sources:
//   IBP_LIMB_FULLRES_HWC  the caller's (H,W,C) map, nearest-pixel lookup exactly like the PAF macro (:9, :322);
//   IBP_LIMB_FEAT_PLANAR  the feature-resolution plane of this limb staged in shared memory (h*w*4 bytes, one
//                         contiguous run, bulk-copy engine); each sample evaluates OpenCV's bicubic x`up` at that
//                         pixel, which equals indexing cv2.resize(pafs, fx=up, fy=up, INTER_CUBIC) (evaluate.py:77-80)
//                         bit for bit without ever writing the 31 MB up-sampled map;
//   ... with twin_base >= 0  the same, staged straight from the raw network output: the upright plane and the
//                         mirrored, channel-permuted plane of the flipped image are averaged on the way into shared
//                         memory (utils/parse_skeletons.py:91-92), so the fused limb maps never exist in HBM.
//
// Scoring is split so that all 256 threads stay busy whatever the number of pairs: per chunk of kPairChunk pairs,
// (0) one thread per pair derives the step geometry, (1) one thread per (pair, sample) evaluates the limb value,
// (2) one thread per pair adds its samples in order (the reference's sequential fp32 sum) and applies the criteria.
//
// Ordering: candidates are compacted in (a-major, b-minor) pair order, then sorted by a bitonic network on
// (score descending, candidate index descending).  That equals libstdc++'s std::sort with the reference's non-strict
// `>=` comparator for n <= 16 (insertion sort: ties end up in reverse input order) and for any n without duplicate
// keys.  A limb with n > 16 AND duplicate keys is re-ordered by one thread running the introsort emulation of
// stl_sort_emul.h (status bit IBP_ST_TIE_EMULATED; IBP_ST_TIE_UNDEFINED when the real sort would have left the array).
#pragma once
#include "ibp_device.cuh"
#include "k_group.cuh"
#include "stl_sort_emul.h"

namespace ibp {

struct LimbArgs {
    const DevPeak *peaks;   // (batch, 18, cap)
    const int *part_count;  // (batch, 18)
    int cap;
    const void *src;        // limb values, see mode
    int dtype;              // element type of src (IBP_F16 only with IBP_LIMB_FEAT_PLANAR)
    int mode;
    size_t image_stride;    // elements between images in src
    int twin_base;          // IBP_LIMB_FEAT_PLANAR: channel offset of the flipped image's block, or -1
    int H, W, C;            // mode 0: full-resolution dims and channel count (stride); sample clamp bounds
    int h, w;               // mode 1: feature dims
    int up;
    int stage_plane;        // mode 1: plane fits the shared buffer
    int min_img_size;
    float thr, step_ratio, w_limb, w_a, w_b;
    int max_steps;
    int cand_cap;           // power of two
    const ibp_peak *line;   // (batch, line_stride) part-major peak list, indexed BY PEAK ID like peak_infos_line
    int line_stride;
    DevLink *links;         // (batch, 30, cap)
    int *link_count;        // (batch, 30)
    uint32_t *status;       // (batch)
    unsigned *ticket;       // (batch) zero on entry, left zero on exit
    int smem_bytes;         // dynamic shared memory of the launch
    int debug;              // experiments only (IBP_DEBUG): 1 = skip grouping, 2 = skip scoring, 32 = publish a limb at once
    const int *order;       // (batch) image handled by grid row y, or null: identity
    // Band-list mode (band_peaks != null): the peak kernel left its per-band lists unpacked and index_bands_kernel
    // counted them (part_count, partbase, order).  Every CTA gathers the lists of its two joint types from the bands
    // with their running ids; the CTAs of limbs 0..17 also write the packed per-part arrays the results are fetched
    // from.
    const DevPeak *band_peaks;  // (batch, 18, nbands, cap)
    const int *band_count;      // (batch, 18, nbands)
    int nbands;
    // streaming kernel only: items (image rank, limb) in heaviest-first order with their image and list sizes
    const int4 *item_meta;      // (batch*31) work queue, see IndexArgs::item_meta
    unsigned *work_counter;     // items beyond the first gridDim.x are claimed here; never reset ...
    unsigned work_base;         // ... its value at launch (a launch adds exactly batch*30)
    int plane_bytes;            // streaming kernel: bytes reserved for the plane at the start of shared memory
    const int *partbase;        // (batch, 19) first id of each part, [18] = peaks of the image (index_bands_kernel)
    int batch;
    DevPeak *peaks_out;         // (batch, 18, cap)
    ibp_peak *line_out;         // (batch, line_stride)
    unsigned long long *timeline;  // experiments only: (batch, 31, 4) %globaltimer stamps, or null
    unsigned char *cta_scratch;    // limbs_stream3_kernel: per-CTA global scratch (cta_scratch_bytes each) for what does not
    size_t cta_scratch_bytes;      // fit its small shared memory: the candidates of a crowded limb, the introsort emulation
    GroupArgs group;
};

constexpr int kLimbThreadsMax = 512;  // CTA size is a template parameter: 256 (default) or 512 (handles sized for crowds)
constexpr int kPairChunk = 128;  // pairs scored per round on the two-phase path
constexpr int kLimbRing = 4;      // slots of the twin-plane ring while staging
constexpr int kMaxDeferBands = 16; // band-list mode: most bands per plane the limb kernel gathers from
constexpr int kRankSortMax = 512; // up to this many candidates are ordered by a rank sort, more by a bitonic network

__device__ __forceinline__ uint32_t float_orderable(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ uint32_t float_orderable_inv(uint32_t k) { return (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; }

template <typename P = float>
struct SamplerT {
    const float *base;  // mode 0: image base; mode 1 with P = float: plane base (shared or global)
    const P *plane;     // mode 1: plane base when P is not float (the streaming kernel's float16 plane)
    const float (*tab4)[4];
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
        if constexpr (std::is_same<P, float>::value) {
            if (up == 4) return cubic_sample4(base, h, w, tab4, px, py);
            return cubic_sample(base, h, w, up, px, py);
        } else {
            if (up == 4) return cubic_sample4(plane, h, w, tab4, px, py);
            return cubic_sample(plane, h, w, up, px, py);
        }
    }
};
using Sampler = SamplerT<float>;

__device__ __forceinline__ float limb_length(const DevPeak &A, const DevPeak &B) {
    const float dx = __int2float_rn(B.x - A.x), dy = __int2float_rn(B.y - A.y);  // :68-69
    return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));        // :70
}

struct PairGeo {  // step geometry of one pair (pafprocess.cpp:66-73, :311-316)
    float ax, ay, sx, sy;
    int steps;  // 0: skipped (zero length)
};

__device__ __forceinline__ PairGeo pair_geometry(const DevPeak &A, const DevPeak &B, int max_steps) {
    const float len = limb_length(A, B);
    PairGeo g;
    g.steps = 0;
    g.ax = __int2float_rn(A.x);
    g.ay = __int2float_rn(A.y);
    g.sx = g.sy = 0.f;
    if (!((double)len < 1e-12)) {                                                                     // :71
        g.steps = min(__double2int_rz(__dadd_rn((double)__fadd_rn(len, 1.0f), 0.5)), max_steps);  // :73
        const float denom = __int2float_rn(g.steps - 1);
        g.sx = __fdiv_rn(__int2float_rn(B.x - A.x), denom);  // :315
        g.sy = __fdiv_rn(__int2float_rn(B.y - A.y), denom);  // :316
    }
    return g;
}
// sample point i of a pair: :319-320 with round2int (:329-331): float multiply, float add, +0.5 in double, truncate
__device__ __forceinline__ void pair_sample_xy(const PairGeo &g, int i, int &px, int &py) {
    const float fi = __int2float_rn(i);
    px = __double2int_rz(__dadd_rn((double)__fadd_rn(g.ax, __fmul_rn(fi, g.sx)), 0.5));
    py = __double2int_rz(__dadd_rn((double)__fadd_rn(g.ay, __fmul_rn(fi, g.sy)), 0.5));
}


// Shared-memory views of one limb item (see the carve-up in the kernels).
struct LimbSmem {
    uint64_t *emu;      // introsort emulation array (cand_cap entries); may alias the plane in the one-shot kernel
    DevPeak *sA, *sB;
    unsigned char *used_a, *used_b;
    float *line_a, *line_b;
    uint64_t *keys;
    float *cand_limb;
    uint32_t *cand_pair;
    PairGeo *geo;
    float *vals;
    int *wcnt;
};

__device__ __forceinline__ void limb_stamp(const LimbArgs &a, int img, int row, int col) {
    if (a.timeline && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.timeline[((size_t)img * 31 + row) * 4 + col] = t;
    }
}

// Orders the `ncand` accepted candidates of one limb (compacted in pair order in m.keys / m.cand_limb / m.cand_pair) and
// connects them greedily (pafprocess.cpp:98-129).  All threads of the CTA; the number of connections written is returned
// on thread 0.  The arrays may live in shared or in global memory.
template <int kLimbThreads, int kRankMax = kRankSortMax>  // kRankMax: the caller's bound on ncand when it is below kRankSortMax
__device__ __forceinline__ int order_and_link(const LimbArgs &a, const LimbSmem &m, int img, int limb, int nA, int nB, int ncand) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t *emu = m.emu;
    DevPeak *sA = m.sA, *sB = m.sB;
    unsigned char *used_a = m.used_a, *used_b = m.used_b;
    float *line_a = m.line_a, *line_b = m.line_b;
    uint64_t *keys = m.keys;
    float *cand_limb = m.cand_limb;
    uint32_t *cand_pair = m.cand_pair;
    int n_links = 0;
    if (ncand > a.cand_cap) {
        if (tid == 0) atomicOr(a.status + img, IBP_ST_CAND_OVERFLOW);
        ncand = a.cand_cap;
    }

    if (ncand > 0) {
        // ---- order: descending on (score, index); keys are distinct because they carry the index ----
        if (ncand <= kRankMax) {
            // rank sort: one pass, two barriers — the common case (a handful of candidates per limb)
            uint64_t mine[kRankMax / kLimbThreads];
            int rank[kRankMax / kLimbThreads];
#pragma unroll
            for (int q = 0; q < kRankMax / kLimbThreads; q++) {
                const int i = tid + q * kLimbThreads;
                mine[q] = (i < ncand) ? keys[i] : 0;
                rank[q] = 0;
            }
            for (int j = 0; j < ncand; j++) {
                const uint64_t kj = keys[j];
#pragma unroll
                for (int q = 0; q < kRankMax / kLimbThreads; q++) rank[q] += (kj > mine[q]);
            }
            __syncthreads();
#pragma unroll
            for (int q = 0; q < kRankMax / kLimbThreads; q++)
                if (tid + q * kLimbThreads < ncand) keys[rank[q]] = mine[q];
            __syncthreads();
        } else {
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

        // ---- greedy connection (:111-129), warp 0: 32 candidates per step, leaders taken in order.  Every lane prepares the
        //      record of its own candidate first (in parallel); the serial part is one ballot and two shuffles per connection:
        //      the leader's two peaks are broadcast and each lane strikes itself out if it shares one of them ----
        if (warp == 0) {
            const int max_links = min(nA, nB);
            DevLink *out = a.links + ((size_t)img * kLimbs + limb) * a.cap;
            int nl = 0;
            for (int base = 0; base < ncand && nl < max_links; base += 32) {
                const int i = base + lane;
                bool valid = i < ncand;
                int ia = -1, ib = -1;  // never equal to a leader's
                DevLink rec;
                rec.id_a = rec.id_b = 0;
                rec.limb = rec.len = rec.gate = rec.pad = rec.score_a = rec.score_b = 0.f;
                if (valid) {
                    const int cidx = (int)(uint32_t)order[i];
                    const uint32_t pr = cand_pair[cidx];
                    const int ca = (int)(pr >> 16), cb = (int)(pr & 0xffffu);
                    valid = !used_a[ca] && !used_b[cb];  // taken in an earlier step of 32
                    if (valid) {
                        ia = ca;
                        ib = cb;
                        const DevPeak A = sA[ia], B = sB[ib];
                        rec.id_a = A.id;
                        rec.id_b = B.id;
                        rec.limb = cand_limb[cidx];
                        rec.len = limb_length(A, B);
                        rec.gate = length_gate(rec.len, a.group.len_rate);
                        rec.score_a = line_a[ia];
                        rec.score_b = line_b[ib];
                    }
                }
                while (nl < max_links) {
                    const uint32_t bal = __ballot_sync(0xffffffffu, valid);
                    if (bal == 0) break;
                    const int leader = __ffs(bal) - 1;
                    const int la = __shfl_sync(0xffffffffu, ia, leader), lb = __shfl_sync(0xffffffffu, ib, leader);
                    if (lane == leader) {
                        used_a[ia] = 1;
                        used_b[ib] = 1;
                        out[nl] = rec;
                    }
                    valid = valid && ia != la && ib != lb;  // the leader itself included
                    nl++;
                }
                __syncwarp();  // the flags, for the next 32
            }
            n_links = nl;
        }
    }
    return n_links;
}

// Scores every (a, b) pair of one limb from the staged plane, orders the accepted candidates and connects them
// greedily (pafprocess.cpp:52-130).  All threads of the CTA; the number of connections written is returned on thread 0.
// `plane_dead` is called by every thread once the last sample has been read (after a barrier): the streaming kernel starts
// the next item's plane copy there, under the ordering / greedy tail.
// kChunk: pairs per round of the two-phase path (m.geo / m.vals hold that many); kRankMax: see order_and_link.
template <int kLimbThreads, int kChunk = kPairChunk, int kRankMax = kRankSortMax, class Smp, class PlaneDead>
__device__ __forceinline__ int score_and_link(const LimbArgs &a, const LimbSmem &m, const Smp &smp, int img, int limb, int nA, int nB,
                                              PlaneDead plane_dead) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    DevPeak *sA = m.sA, *sB = m.sB;
    uint64_t *keys = m.keys;
    float *cand_limb = m.cand_limb;
    uint32_t *cand_pair = m.cand_pair;
    PairGeo *geo = m.geo;
    float *vals = m.vals;
    int *wcnt = m.wcnt;
    // ---- score all pairs; compact the accepted ones in pair order ----
    const int npairs = (a.debug & 2) ? 0 : nA * nB;
    int ncand = 0;
    // Few pairs (the usual image): split every pair's samples over the threads (phases 0-2 below).  Many pairs (a
    // crowd): one thread per pair does its own samples — enough parallelism already, no staging, fewer barriers.
    const bool direct = npairs > kLimbThreads;
    const int chunk_size = direct ? kLimbThreads : kChunk;
    const bool fastdiv = a.max_steps <= 32;
    const unsigned rcp_steps = (1u << 20) / (unsigned)a.max_steps + 1u;
    const double half_size = __dmul_rn(0.5, (double)a.min_img_size);
    for (int base = 0; base < npairs; base += chunk_size) {
        const int chunk = min(chunk_size, npairs - base);
        if (!direct) {
            // (0) geometry, one thread per pair
            if (tid < chunk) {
                const int p = base + tid;
                const int ia = p / nB, ib = p - ia * nB;
                geo[tid] = pair_geometry(sA[ia], sB[ib], a.max_steps);
            }
            __syncthreads();
            // (1) samples, one thread per (pair, step)
            for (int it = tid; it < chunk * a.max_steps; it += kLimbThreads) {
                // it / max_steps: exact by multiply and shift for it < 4096 and max_steps <= 32 (error < 2^-8 < 1 / max_steps)
                const int q = fastdiv ? (int)(((unsigned)it * rcp_steps) >> 20) : it / a.max_steps, i = it - q * a.max_steps;
                const PairGeo g = geo[q];
                if (i < g.steps) {
                    int px, py;
                    pair_sample_xy(g, i, px, py);
                    vals[it] = smp.at(px, py);
                }
            }
            __syncthreads();
        }
        // (2) criteria, one thread per pair
        bool ok = false;
        float ls = 0.f, rk = 0.f;
        int ia = 0, ib = 0;
        if (tid < chunk) {
            const int p = base + tid;
            ia = p / nB;
            ib = p - ia * nB;
            const DevPeak A = sA[ia], B = sB[ib];
            const PairGeo g = direct ? pair_geometry(A, B, a.max_steps) : geo[tid];
            if (g.steps > 0) {
                const float len = limb_length(A, B);
                const float *v = vals + tid * a.max_steps;
                float sum = 0.0f;
                int above = 0;
                for (int i = 0; i < g.steps; i++) {
                    float vi;
                    if (direct) {
                        int px, py;
                        pair_sample_xy(g, i, px, py);
                        vi = smp.at(px, py);
                    } else {
                        vi = v[i];
                    }
                    sum = __fadd_rn(sum, vi);   // :85
                    above += (vi > a.thr);      // :87
                }
                // :90  float + min(0.0, 0.5*min_img_size/len - 1.0) evaluated in double, narrowed on assignment.  A limb shorter
                // than half the image size pays no penalty: the exact quotient is above 1, so is the rounded one, the min is 0
                // and the sum of a float and 0.0 narrows back to that float — no double division for the usual limb
                ls = __fdiv_rn(sum, __int2float_rn(g.steps));
                if (!((double)len < half_size)) {
                    const double pen_raw = __dsub_rn(__ddiv_rn(half_size, (double)len), 1.0);
                    const double pen = (pen_raw < 0.0) ? pen_raw : 0.0;
                    ls = __double2float_rn(__dadd_rn((double)ls, pen));
                }
                const float need = __fmul_rn(__int2float_rn(g.steps), a.step_ratio);  // :91
                if (__int2float_rn(above) > need && ls > 0.0f) {                      // :93
                    ok = true;
                    rk = __fadd_rn(__fadd_rn(__fmul_rn(a.w_limb, ls), __fmul_rn(a.w_a, A.score)), __fmul_rn(a.w_b, B.score));  // :94-96
                }
            }
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
    plane_dead();
    limb_stamp(a, img, limb, 2);
    return order_and_link<kLimbThreads, kRankMax>(a, m, img, limb, nA, nB, ncand);
}

template <typename T, int kLimbThreads>
__global__ void __launch_bounds__(kLimbThreads) limbs_kernel(LimbArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int limb = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int pa = c_limb_a[limb], pb = c_limb_b[limb];
    const bool from_bands = a.band_peaks != nullptr;
    __shared__ int s_partbase[kParts + 1];  // band-list mode: first id of each part; [18] = peaks of the image
    __shared__ int s_bc[kParts * kMaxDeferBands];  // band-list mode: the image's band counts
    const int img = a.order ? a.order[blockIdx.y] : (int)blockIdx.y;
    const int nA = a.part_count[img * kParts + pa], nB = a.part_count[img * kParts + pb];
    if (from_bands) {  // published by the barrier that follows the start of the plane copies
        const int items = kParts * a.nbands;
        for (int i = tid; i < items; i += kLimbThreads) s_bc[i] = a.band_count[(size_t)img * items + i];
        if (tid <= kParts) s_partbase[tid] = a.partbase[img * (kParts + 1) + tid];
    }
    int *link_count = a.link_count + img * kLimbs + limb;
    // band-list mode: k-th peak of part p (k < its count) = entry (band, k - first of band); copies it with its id
    auto gather_part = [&](int p, int k, DevPeak &out) {
        const int *bc = s_bc + p * a.nbands;
        int bnd = 0, first = 0;
        for (; bnd + 1 < a.nbands; bnd++) {
            const int c = bc[bnd];
            if (k < first + c) break;
            first += c;
        }
        out = a.band_peaks[(((size_t)img * kParts + p) * a.nbands + bnd) * a.cap + (k - first)];
        out.id = s_partbase[p] + k;
    };

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
    DevPeak *sA = reinterpret_cast<DevPeak *>(smem_raw + off);
    off += (size_t)a.cap * sizeof(DevPeak);
    DevPeak *sB = reinterpret_cast<DevPeak *>(smem_raw + off);
    off += (size_t)a.cap * sizeof(DevPeak);
    unsigned char *used_a = smem_raw + off;
    off += (size_t)a.cap;
    unsigned char *used_b = smem_raw + off;
    off += (size_t)a.cap;
    off = (off + 15) & ~(size_t)15;
    float *line_a = reinterpret_cast<float *>(smem_raw + off);  // peak_infos_line[id].score of each candidate end
    off += (size_t)a.cap * 4;
    float *line_b = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)a.cap * 4;
    off = (off + 15) & ~(size_t)15;
    // everything from here on is idle while the plane is staged: the twin ring borrows it
    const size_t ring_off = off;
    uint64_t *keys = reinterpret_cast<uint64_t *>(smem_raw + off);
    off += (size_t)a.cand_cap * 8;
    float *cand_limb = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)a.cand_cap * 4;
    uint32_t *cand_pair = reinterpret_cast<uint32_t *>(smem_raw + off);
    off += (size_t)a.cand_cap * 4;
    PairGeo *geo = reinterpret_cast<PairGeo *>(smem_raw + off);  // [kPairChunk]
    off += (size_t)kPairChunk * sizeof(PairGeo);
    float *vals = reinterpret_cast<float *>(smem_raw + off);     // [kPairChunk][max_steps]
    off += (size_t)kPairChunk * a.max_steps * 4;
    const int ring_bytes = (int)((off - ring_off) & ~(size_t)15);
    __shared__ uint64_t bars[1 + kLimbRing];
    __shared__ int wcnt[kLimbThreads / 32];
    __shared__ int s_last;
    __shared__ float tab4[4][4];

    int n_links = 0;  // meaningful on thread 0 only
    auto stamp = [&](int row, int col) {
        if (a.timeline && tid == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.timeline[((size_t)img * 31 + row) * 4 + col] = t;
        }
    };
    // band-list mode: the CTA of limb l < 18 also writes part l of the packed arrays the results are fetched from
    // (evaluate.py:99-108 joint list); called while the plane copies are in flight
    auto pack_part = [&]() {
        if (!from_bands || limb >= kParts) return;
        const int c = s_partbase[limb + 1] - s_partbase[limb];
        const int n = min(c, a.cap);
        for (int k = tid; k < n; k += kLimbThreads) {
            DevPeak p;
            gather_part(limb, k, p);
            a.peaks_out[((size_t)img * kParts + limb) * a.cap + k] = p;
            if (p.id < a.line_stride) {
                ibp_peak o;
                o.x = p.x;
                o.y = p.y;
                o.score = p.score;
                o.part = limb;
                a.line_out[(size_t)img * a.line_stride + p.id] = o;
            }
        }
    };
    stamp(limb, 0);
    // no pair => no candidate => no connection (covers the both-empty skip at :57)
    if (!(nA > 0 && nB > 0) && from_bands) {
        __syncthreads();
        pack_part();
    }
    if (nA > 0 && nB > 0) {
        // source planes of this limb (feature-resolution modes)
        const T *gplane = nullptr, *twin = nullptr;
        if (a.mode == IBP_LIMB_FEAT_PLANAR) {
            const T *base = reinterpret_cast<const T *>(a.src) + (size_t)img * a.image_stride;
            const size_t hw = (size_t)a.h * a.w;
            gplane = base + (size_t)limb * hw;
            // flipped image: channel flip_paf_ord[limb] of its block (config/config.py:150-152)
            twin = (a.twin_base >= 0) ? base + (size_t)(a.twin_base + c_flip_src[limb]) * hw : nullptr;
        }
        // twin rows stream through a ring that borrows the candidate/geometry/sample scratch (idle until scoring)
        float *ring = reinterpret_cast<float *>(smem_raw + ring_off);
        const int slot_rows = (a.mode == IBP_LIMB_FEAT_PLANAR) ? min(a.h, (ring_bytes / kLimbRing) / (a.w * 4)) : 0;
        bool tma_flip = false;
        if (sizeof(T) == 4 && a.mode == IBP_LIMB_FEAT_PLANAR && a.stage_plane)
            tma_flip = flip_tma_ok(plane_s, reinterpret_cast<const float *>(gplane), reinterpret_cast<const float *>(twin), ring, a.w, slot_rows);
        if (tid == 0) {
            for (int i = 0; i <= kLimbRing; i++) mbar_init(&bars[i], 1);
            fence_barrier_init();
            // start the copies before anything else so that they overlap the candidate-list loads below
            if (tma_flip)
                flip_tma_issue<kLimbRing>(plane_s, reinterpret_cast<const float *>(gplane), reinterpret_cast<const float *>(twin), 0, a.h, a.w, ring,
                               slot_rows, bars);
        }
        cubic_tab4_init(tab4);
        const DevPeak *gA = a.peaks + ((size_t)img * kParts + pa) * a.cap;
        const DevPeak *gB = a.peaks + ((size_t)img * kParts + pb) * a.cap;
        // pafprocess.cpp:160,261-262 look the peak scores up as peak_infos_line[peak_id]: the part-major list indexed
        // by ARRIVAL id — the same peak only when the caller's joint list is sorted by part (evaluate.py:99-103 does
        // that); keep the quirk, but fetch the values now so that the greedy loop never waits on global memory
        const ibp_peak *line = a.line + (size_t)img * a.line_stride;
        if (from_bands) {
            __syncthreads();  // s_bc / s_partbase
            for (int i = tid; i < nA; i += kLimbThreads) {
                DevPeak p;
                gather_part(pa, i, p);
                sA[i] = p;
                used_a[i] = 0;
                line_a[i] = p.score;  // ids are arrival order here, so peak_infos_line[id] is the peak itself
            }
            for (int i = tid; i < nB; i += kLimbThreads) {
                DevPeak p;
                gather_part(pb, i, p);
                sB[i] = p;
                used_b[i] = 0;
                line_b[i] = p.score;
            }
        } else {
        for (int i = tid; i < nA; i += kLimbThreads) {
            const DevPeak p = gA[i];
            sA[i] = p;
            used_a[i] = 0;
            line_a[i] = line[min(max(p.id, 0), a.line_stride - 1)].score;
        }
        for (int i = tid; i < nB; i += kLimbThreads) {
            const DevPeak p = gB[i];
            sB[i] = p;
            used_b[i] = 0;
            line_b[i] = line[min(max(p.id, 0), a.line_stride - 1)].score;
        }
        }
        pack_part();
        __syncthreads();

        Sampler smp;
        smp.plane = nullptr;
        smp.mode = a.mode;
        smp.limb = limb;
        smp.H = a.H;
        smp.W = a.W;
        smp.C = a.C;
        smp.h = a.h;
        smp.w = a.w;
        smp.up = a.up;
        smp.tab4 = tab4;
        if (a.mode == IBP_LIMB_FEAT_PLANAR) {
            if (a.stage_plane) {
                if (tma_flip) {
                    flip_tma_finish<kLimbRing>(plane_s, reinterpret_cast<const float *>(twin), 0, a.h, a.w, ring, slot_rows, bars);
                } else {
                    uint32_t phase = 0;
                    stage_rows_fused<T>(plane_s, gplane, twin, 0, a.h, a.w, &bars[0], phase);
                }
                smp.base = plane_s;
            } else {
                smp.base = reinterpret_cast<const float *>(gplane);  // host side guarantees float + no twin here
            }
        } else {
            smp.base = reinterpret_cast<const float *>(a.src) + (size_t)img * a.image_stride;
        }

        stamp(limb, 1);
        LimbSmem m;
        m.emu = emu;
        m.sA = sA;
        m.sB = sB;
        m.used_a = used_a;
        m.used_b = used_b;
        m.line_a = line_a;
        m.line_b = line_b;
        m.keys = keys;
        m.cand_limb = cand_limb;
        m.cand_pair = cand_pair;
        m.geo = geo;
        m.vals = vals;
        m.wcnt = wcnt;
        n_links = score_and_link<kLimbThreads>(a, m, smp, img, limb, nA, nB, [] {});
    }
    if (tid == 0) *link_count = n_links;
    stamp(limb, 3);

    // ---- the last CTA of this image groups its connections into people ----
    if (warp == 0) __threadfence();  // only warp 0 wrote connections / the count
    __syncthreads();
    if (tid == 0) {
        const unsigned t = atomicAdd(a.ticket + img, 1u);
        s_last = (t == (unsigned)kLimbs - 1);
        if (s_last) a.ticket[img] = 0;  // ready for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (a.debug & 1) return;
    stamp(30, 0);
    group_image(a.group, img, smem_raw, (size_t)a.smem_bytes);
    stamp(30, 1);
}

// ---------------------------------------------------------------------------------------------------------------
// Streaming variant: a persistent grid (two CTAs per SM) walks one work queue — per image, heaviest image first: its
// 30 limbs, then "group this image" — item blockIdx.x first, then whatever the shared counter hands out.  Only for
// the pipeline's common case: float32 network output with the flip twin, plane staged in shared memory, peak lists
// left in bands by the streaming peak kernel (item_meta / partbase from index_bands_kernel).
// The point of the loop: the plane buffer is dead once the last sample of an item has been read, long before the item
// is finished (ordering, greedy selection, connection records, ticket, possibly the grouping of the whole image).
// Thread 0 claims the next item at the top of each round, learns its image and list sizes while the CTA stages and
// scores, and starts the bulk copy of the NEXT item's upright plane the moment scoring ends — so that plane is
// (mostly) there when the next round begins and the round only streams the twin rows through the ring.
// Grouping is a queue item of its own so that no CTA ever sits on a claimed limb while it runs a long chain: whoever
// claims it waits (holding nothing else) until the image's 30 limbs have published their connections — they were
// claimed just before by CTAs that never wait for anybody, so the wait is bounded by one limb's run time — groups, and
// only then claims again.  The grouping tables live behind the plane buffer, so a copy in flight is never in their way.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kLimbStreamThreads = 256;
constexpr int kLimbStreamThreadsHalf = 128;  // float16 planes, see limbs_stream_kernel

// BOX: stage only the columns an item's samples can touch (the source is host memory read over the bus); from device memory
// whole rows are staged — one bulk copy per item beats a copy per row there.
// THREADS: 256 for float32 planes (64 KB: two CTAs per SM either way); 128 for float16 planes, whose 32 KB let a third CTA
// in — an item is mostly a serial chain, four warps run it about as fast as eight, and three items in flight beat two
// (measured on configs[1] in float16: 765 k img/s against 649 k with 256 threads).
template <typename T, bool BOX = false, int THREADS = kLimbStreamThreads>
__global__ void __launch_bounds__(THREADS, 2) limbs_stream_kernel(LimbArgs a) {
    constexpr int kLimbThreads = THREADS;
    constexpr int V = 16 / (int)sizeof(T);  // elements per 16-byte vector
    using Vec = typename Vec16<T>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int items = a.batch * (kLimbs + 1);
    const size_t hw = (size_t)a.h * a.w;

    // ---- carve shared memory: plane | peak lists | ring = candidate / geometry / sample scratch ----
    T *plane_s = reinterpret_cast<T *>(smem_raw);
    size_t off = (size_t)a.plane_bytes;
    LimbSmem m;
    m.sA = reinterpret_cast<DevPeak *>(smem_raw + off);
    off += (size_t)a.cap * sizeof(DevPeak);
    m.sB = reinterpret_cast<DevPeak *>(smem_raw + off);
    off += (size_t)a.cap * sizeof(DevPeak);
    m.used_a = smem_raw + off;
    off += (size_t)a.cap;
    m.used_b = smem_raw + off;
    off += (size_t)a.cap;
    off = (off + 15) & ~(size_t)15;
    m.line_a = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)a.cap * 4;
    m.line_b = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)a.cap * 4;
    off = (off + 15) & ~(size_t)15;
    const size_t ring_off = off;
    m.keys = reinterpret_cast<uint64_t *>(smem_raw + off);
    off += (size_t)a.cand_cap * 8;
    m.cand_limb = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)a.cand_cap * 4;
    m.cand_pair = reinterpret_cast<uint32_t *>(smem_raw + off);
    off += (size_t)a.cand_cap * 4;
    m.geo = reinterpret_cast<PairGeo *>(smem_raw + off);
    m.emu = reinterpret_cast<uint64_t *>(smem_raw + off);  // geometry + sample scratch are idle by then (host checks the size)
    off += (size_t)kPairChunk * sizeof(PairGeo);
    m.vals = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)kPairChunk * a.max_steps * 4;
    const int ring_bytes = (int)((off - ring_off) & ~(size_t)15);
    T *ring = reinterpret_cast<T *>(smem_raw + ring_off);
    const int slot_rows = min(a.h, (ring_bytes / kLimbRing) / (a.w * (int)sizeof(T)));
    __shared__ uint64_t bars[1 + kLimbRing];
    __shared__ int wcnt[kLimbThreads / 32];
    __shared__ int s_last;
    __shared__ float tab4[4][4];
    __shared__ int s_partbase[kParts + 1];
    __shared__ int s_bc[kParts * kMaxDeferBands];
    __shared__ int4 s_next;  // {image, nA | nB << 16, limb (30: group the image), r0 | r1 << 16}; image < 0: no more work
    m.wcnt = wcnt;

    auto claim_to_item = [&](unsigned t) {
        const unsigned rel = t - a.work_base + gridDim.x;
        return rel < (unsigned)items ? (int)rel : -1;
    };
    auto upright_of = [&](int img, int limb) {
        return reinterpret_cast<const T *>(a.src) + (size_t)img * a.image_stride + (size_t)limb * hw;
    };
    auto twin_of = [&](int img, int limb) {  // flipped image: channel flip_paf_ord[limb] of its block (config/config.py:150-152)
        return reinterpret_cast<const T *>(a.src) + (size_t)img * a.image_stride + (size_t)(a.twin_base + c_flip_src[limb]) * hw;
    };
    // Only the box of feature rows [r0, r1) x columns [c0, c1) that a sample of the item can touch is staged (at its own place in
    // the buffer): item.w = r0 | r1 << 16, item.z = limb | (c0 / 4) << 8 | (c1 / 4) << 20 (index_bands_kernel).
    auto item_limb = [](const int4 &it) { return it.z & 0xff; };
    auto item_c0 = [&](const int4 &it) { return BOX ? ((it.z >> 8) & 0xfff) << 2 : 0; };
    auto item_c1 = [&](const int4 &it) { return BOX ? (int)((unsigned)it.z >> 20) << 2 : a.w; };
    auto has_pairs = [&](const int4 &it) { return (it.y & 0xffff) > 0 && (it.y >> 16) > 0 && item_limb(it) < kLimbs; };
    // one thread: `rows` rows of columns [col0, col0 + ncols) from src to dst (both with a row pitch of a.w elements)
    auto bulk_box = [&](T *dst, const T *src, int rows, int col0, int ncols, uint64_t *bar) {
        if (!BOX || ncols == a.w) {
            bulk_rows(dst, src, (uint32_t)rows * a.w * (uint32_t)sizeof(T), bar);
            return;
        }
        const uint32_t row_bytes = (uint32_t)ncols * (uint32_t)sizeof(T);
        mbar_expect_tx(bar, (uint32_t)rows * row_bytes);
        for (int r = 0; r < rows; r++) bulk_g2s(dst + (size_t)r * a.w + col0, src + (size_t)r * a.w + col0, row_bytes, bar);
    };
    // one thread: chunk c of the twin rows into its ring slot; the twin is mirrored, the box's columns [c0, c1) are its [w - c1, w - c0)
    auto twin_chunk = [&](const T *twin, int r0, int r1, int c0, int c1, int c) {
        const int s = c % kLimbRing, first = r0 + c * slot_rows;
        const int rows = min(slot_rows, r1 - first);
        bulk_box(ring + (size_t)s * slot_rows * a.w, twin + (size_t)first * a.w, rows, a.w - c1, c1 - c0, &bars[1 + s]);
    };
    auto upright_rows = [&](const int4 &it) {  // one thread: the upright box of a limb item straight to its place
        if (!has_pairs(it)) return;
        const int r0 = it.w & 0xffff, r1 = it.w >> 16;
        bulk_box(plane_s + (size_t)r0 * a.w, upright_of(it.x, item_limb(it)) + (size_t)r0 * a.w, r1 - r0, item_c0(it), item_c1(it) - item_c0(it), &bars[0]);
    };

    pdl_wait();  // launched as a programmatic dependent: the work queue and the band lists come from the kernels before
    if (tid == 0) {
        for (int i = 0; i <= kLimbRing; i++) mbar_init(&bars[i], 1);
        fence_barrier_init();
        const int4 first = a.item_meta[blockIdx.x];  // the grid never exceeds the item count
        upright_rows(first);
        s_next = first;
    }
    cubic_tab4_init(tab4);
    __syncthreads();

    uint32_t up_uses = 0;   // completed waits on bars[0] (one per item with pairs)
    uint32_t ring_par = 0;  // bit s: parity of the next wait on ring slot s
    // Publishing a limb (fence, then +1 on the image's ticket) is put off until the stores it orders have had the time of
    // the next item's first memory round trip to reach L2: issued right behind them, the fence holds the whole CTA for
    // that long.  Nothing waits in between: a group item publishes what is pending before it starts to wait.
    int pending_img = -1;
    auto publish_pending = [&]() {
        if (pending_img >= 0 && tid < 32) {
            __threadfence();  // warp 0 wrote the connections and the count
            __syncwarp();
            if (tid == 0) atomicAdd(a.ticket + pending_img, 1u);
        }
        pending_img = -1;
    };
    for (;;) {
        const int4 cur = s_next;
        if (cur.x < 0) break;
        const int img = cur.x, nA = cur.y & 0xffff, nB = cur.y >> 16, limb = item_limb(cur);
        const int r0 = cur.w & 0xffff, r1 = cur.w >> 16, c0 = item_c0(cur), c1 = item_c1(cur);
        unsigned claimed = 0;
        int4 nmeta = make_int4(-1, 0, 0, 0);
        // thread 0: who is next (one load, issued early), and — once the plane is dead — start its upright plane
        auto resolve_next = [&]() {
            if (tid == 0) {
                const int nitem = claim_to_item(claimed);
                if (nitem >= 0) nmeta = a.item_meta[nitem];
            }
        };
        auto prefetch_next = [&]() {
            if (tid == 0) {
                if (nmeta.x >= 0 && has_pairs(nmeta)) {
                    fence_proxy_async();  // the samples read through the generic proxy come first
                    upright_rows(nmeta);
                }
                s_next = nmeta;
            }
        };
        if (limb == kLimbs) {
            // ---- group the image (pafprocess.cpp:133-285) once its 30 limbs are in ----
            publish_pending();
            if (tid == 0) {
                unsigned seen = 0;
                for (int spins = 0; spins < (1 << 22); spins++) {
                    asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(seen) : "l"(a.ticket + img) : "memory");
                    if (seen >= (unsigned)kLimbs) break;
                    __nanosleep(100);
                }
                s_last = seen >= (unsigned)kLimbs;
                if (s_last) a.ticket[img] = 0;  // ready for the next launch
                else atomicOr(a.status + img, IBP_ST_STALLED);
            }
            __syncthreads();
            if (s_last && !(a.debug & 1)) {
                limb_stamp(a, img, 30, 0);
                group_image(a.group, img, smem_raw + a.plane_bytes, (size_t)a.smem_bytes - (size_t)a.plane_bytes,
                            a.timeline ? a.timeline + ((size_t)img * 31 + 30) * 4 + 2 : nullptr);
                limb_stamp(a, img, 30, 1);
            }
            // claim only now: nothing must wait behind a grouping chain
            if (tid == 0) claimed = atomicAdd(a.work_counter, 1u);
            resolve_next();
            prefetch_next();
            __syncthreads();
            continue;
        }
        const int pa = c_limb_a[limb], pb = c_limb_b[limb];
        const bool pairs = nA > 0 && nB > 0;
        const int nchunks = (r1 - r0 + slot_rows - 1) / slot_rows;
        const T *twin = twin_of(img, limb);
        // thread 0: next claim out now, answer used later; the ring is free again: first twin chunks out
        if (tid == 0) {
            claimed = atomicAdd(a.work_counter, 1u);
            if (pairs) {
                fence_proxy_async();  // the previous round's generic-proxy use of the ring region comes first
                for (int c = 0; c < kLimbRing && c < nchunks; c++) twin_chunk(twin, r0, r1, c0, c1, c);
            }
        }
        limb_stamp(a, img, limb, 0);
        {  // band counts / first ids of this image
            const int nitems = kParts * a.nbands;
            for (int i = tid; i < nitems; i += kLimbThreads) s_bc[i] = a.band_count[(size_t)img * nitems + i];
            if (tid <= kParts) s_partbase[tid] = a.partbase[img * (kParts + 1) + tid];
        }
        __syncthreads();
        publish_pending();  // the limb this CTA finished before this one
        // k-th peak of part p = entry (band, k - first of band) of the band lists, with its running id
        auto gather_part = [&](int p, int k, DevPeak &out) {
            const int *bc = s_bc + p * a.nbands;
            int bnd = 0, first = 0;
            for (; bnd + 1 < a.nbands; bnd++) {
                const int c = bc[bnd];
                if (k < first + c) break;
                first += c;
            }
            out = a.band_peaks[(((size_t)img * kParts + p) * a.nbands + bnd) * a.cap + (k - first)];
            out.id = s_partbase[p] + k;
        };
        // the CTA of limb l < 18 writes part l of the packed arrays the results are fetched from (evaluate.py:99-108)
        auto pack_part = [&]() {
            if (limb >= kParts) return;
            const int n = min(s_partbase[limb + 1] - s_partbase[limb], a.cap);
            for (int k = tid; k < n; k += kLimbThreads) {
                DevPeak p;
                gather_part(limb, k, p);
                a.peaks_out[((size_t)img * kParts + limb) * a.cap + k] = p;
                if (p.id < a.line_stride) {
                    ibp_peak o;
                    o.x = p.x;
                    o.y = p.y;
                    o.score = p.score;
                    o.part = limb;
                    a.line_out[(size_t)img * a.line_stride + p.id] = o;
                }
            }
        };
        int n_links = 0;
        if (!pairs) {  // no pair => no candidate => no connection (covers the both-empty skip at :57)
            pack_part();
            resolve_next();
            prefetch_next();
        } else {
            for (int i = tid; i < nA; i += kLimbThreads) {
                DevPeak p;
                gather_part(pa, i, p);
                m.sA[i] = p;
                m.used_a[i] = 0;
                m.line_a[i] = p.score;  // ids are arrival order here, so peak_infos_line[id] is the peak itself
            }
            for (int i = tid; i < nB; i += kLimbThreads) {
                DevPeak p;
                gather_part(pb, i, p);
                m.sB[i] = p;
                m.used_b[i] = 0;
                m.line_b[i] = p.score;
            }
            pack_part();
            resolve_next();
            // ---- fold the twin rows into the upright plane as they land (utils/parse_skeletons.py:91-92) ----
            mbar_wait(&bars[0], up_uses & 1u);
            up_uses++;
            {
                const int wv = a.w / V;
                const int v0 = c0 / V, nv = (c1 - c0) / V;  // the box in 16-byte vectors
                const int shift = ((nv & (nv - 1)) == 0) ? (__ffs(nv) - 1) : -1;
                Vec *d4 = reinterpret_cast<Vec *>(plane_s);
                for (int c = 0; c < nchunks; c++) {
                    const int s = c % kLimbRing;
                    mbar_wait(&bars[1 + s], (ring_par >> s) & 1u);
                    ring_par ^= 1u << s;
                    const int first = r0 + c * slot_rows;
                    const int rows = min(slot_rows, r1 - first);
                    const Vec *r4 = reinterpret_cast<const Vec *>(ring + (size_t)s * slot_rows * a.w);
                    Vec *o4 = d4 + (size_t)first * wv;
                    if (!BOX || nv == wv) {  // whole rows
                        for (int i = tid; i < rows * wv; i += kLimbThreads) {
                            const int r = (shift >= 0) ? (i >> shift) : (i / wv);
                            const int col = i - r * wv;
                            Vec v = o4[i];
                            fold_mirrored(v, r4[r * wv + (wv - 1 - col)]);
                            o4[i] = v;
                        }
                    } else {
                        for (int i = tid; i < rows * nv; i += kLimbThreads) {
                            const int r = (shift >= 0) ? (i >> shift) : (i / nv);
                            const int col = v0 + i - r * nv;
                            Vec v = o4[r * wv + col];
                            fold_mirrored(v, r4[r * wv + (wv - 1 - col)]);
                            o4[r * wv + col] = v;
                        }
                    }
                    if (c + kLimbRing < nchunks) {
                        __syncthreads();  // everybody has read slot s
                        if (tid == 0) {
                            fence_proxy_async();
                            twin_chunk(twin, r0, r1, c0, c1, c + kLimbRing);
                        }
                    }
                }
                __syncthreads();
            }
            limb_stamp(a, img, limb, 1);
            SamplerT<T> smp;
            smp.mode = IBP_LIMB_FEAT_PLANAR;
            smp.limb = limb;
            smp.H = a.H;
            smp.W = a.W;
            smp.C = a.C;
            smp.h = a.h;
            smp.w = a.w;
            smp.up = a.up;
            smp.tab4 = tab4;
            smp.base = reinterpret_cast<const float *>(plane_s);
            smp.plane = plane_s;
            n_links = score_and_link<kLimbThreads>(a, m, smp, img, limb, nA, nB, prefetch_next);
        }
        if (tid == 0) a.link_count[img * kLimbs + limb] = n_links;
        limb_stamp(a, img, limb, 3);

        // ---- connections and count of this limb are out (warp 0 wrote them); the image's group item counts 30 tickets ----
        pending_img = img;
        if (a.debug & 32) publish_pending();  // experiments: publish at once
        __syncthreads();  // s_next, and every use of this round's shared data
    }
    publish_pending();
}

}  // namespace ibp
