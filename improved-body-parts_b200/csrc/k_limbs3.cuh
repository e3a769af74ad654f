// K3 + K4, the streaming limb kernel with THREE CTAs per SM (device-resident network output).
//
// Same work queue, same arithmetic and the same publish / grouping protocol as limbs_stream_kernel (k_limbs.cuh); what
// changes is how much shared memory an item in flight holds.  A limb item is a chain of dependent latencies (claim ->
// lists -> plane -> fold -> score -> order -> publish) with little arithmetic, so the kernel's rate is items in flight x
// 1 / chain length; limbs_stream_kernel keeps two items per SM because each carries 36 KB of scratch next to its 64 KB
// plane.  Here the scratch is 9.3 KB, three CTAs fit an SM, and the chain loses its CTA-wide barriers in the middle:
//   * the twin rows do not pass through a shared-memory ring: every thread loads 16-byte pieces of the mirrored twin
//     rows straight into registers (two batches in flight, the first issued before the wait for the upright plane) and
//     folds them into the plane the bulk copy delivered;
//   * pairs are scored by warps, not by the CTA: lane q of a warp owns pair 8q + warp (geometry and criteria, one pair
//     per lane), and the warp walks its pairs with lane i evaluating sample i — the samples of a pair are added in the
//     reference's order through shuffles (pafprocess.cpp:85), no staging of values in shared memory, no barrier;
//   * accepted pairs set a bit of a 256-bit map; a candidate's place in pair order is the number of bits below its own;
//   * the rare cases that need room — more than 256 pairs (a crowd), the introsort emulation — use a per-CTA scratch in
//     global memory; the grouping tables use the plane's space (no copy is in flight while an image is grouped).
#pragma once
#include "k_limbs.cuh"

namespace ibp {

constexpr int kL3Threads = 256;
constexpr int kL3Pairs = 256;  // pairs (and so candidates) of an item scored by the warp path
constexpr int kL3FoldU = 4;    // 16-byte twin pieces per thread and batch
constexpr int kL3MaxBands = 8; // bands per plane this kernel keeps the counts of (two sets: the item's and the next item's)
constexpr int kL3WarpVals = kL3Pairs * 4 / (kL3Threads / 32);  // sample values a warp stages at a time (its share of the candidate arrays)

// shared memory of a CTA beyond the plane: two peak lists with their flags and scores, 256 candidates
__host__ __device__ inline size_t l3_scratch_bytes(int cap) {
    size_t off = (size_t)cap * (2 * sizeof(DevPeak) + 2);
    off = (off + 15) & ~(size_t)15;
    off += (size_t)cap * 8;
    off = (off + 15) & ~(size_t)15;
    return off + (size_t)kL3Pairs * 16;
}
// global scratch of a CTA: keys, limb scores, pairs of cand_cap candidates + the emulation array
__host__ __device__ inline size_t l3_global_scratch_bytes(int cand_cap) { return (size_t)cand_cap * 24; }

__device__ __forceinline__ void cp_async4(void *dst_smem, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(kL3Threads, 3) limbs_stream3_kernel(LimbArgs a) {
    constexpr int kLimbThreads = kL3Threads;
    constexpr int V = 16 / (int)sizeof(T);  // elements per 16-byte vector
    using Vec = typename Vec16<T>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int items = a.batch * (kLimbs + 1);
    const size_t hw = (size_t)a.h * a.w;

    // ---- carve shared memory: plane | peak lists | candidates ----
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
    uint64_t *keys_s = reinterpret_cast<uint64_t *>(smem_raw + off);
    float *cand_limb_s = reinterpret_cast<float *>(smem_raw + off + (size_t)kL3Pairs * 8);
    uint32_t *cand_pair_s = reinterpret_cast<uint32_t *>(smem_raw + off + (size_t)kL3Pairs * 12);
    unsigned char *gs = a.cta_scratch + (size_t)blockIdx.x * a.cta_scratch_bytes;
    uint64_t *keys_g = reinterpret_cast<uint64_t *>(gs);
    float *cand_limb_g = reinterpret_cast<float *>(gs + (size_t)a.cand_cap * 8);
    uint32_t *cand_pair_g = reinterpret_cast<uint32_t *>(gs + (size_t)a.cand_cap * 12);
    m.emu = reinterpret_cast<uint64_t *>(gs + (size_t)a.cand_cap * 16);
    m.geo = nullptr;
    m.vals = nullptr;
    __shared__ uint64_t bar;  // the upright plane of the item
    __shared__ int wcnt[kLimbThreads / 32];
    __shared__ int s_last;
    __shared__ float tab4[4][4];
    __shared__ int s_pbase[2][kParts + 1];          // first id of each part of the image, two sets: see s_bcnt
    __shared__ int s_bcnt[2][kParts * kL3MaxBands];  // band counts of the image: the item's set and the one the next item's are copied into
    __shared__ int s_nimg;                           // image of the next item if it is a limb item, else -1
    __shared__ int4 s_next;  // {image, nA | nB << 16, limb (30: group the image), r0 | r1 << 16}; image < 0: no more work
    __shared__ unsigned s_accept[kL3Pairs / 32];
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
    auto item_limb = [](const int4 &it) { return it.z & 0xff; };
    auto has_pairs = [&](const int4 &it) { return (it.y & 0xffff) > 0 && (it.y >> 16) > 0 && item_limb(it) < kLimbs; };
    // one thread: the feature rows [r0, r1) of a limb item's upright plane straight to their place (item.w = r0 | r1 << 16)
    auto upright_rows = [&](const int4 &it) {
        if (!has_pairs(it)) return;
        const int r0 = it.w & 0xffff, r1 = it.w >> 16;
        bulk_rows(plane_s + (size_t)r0 * a.w, upright_of(it.x, item_limb(it)) + (size_t)r0 * a.w, (uint32_t)(r1 - r0) * a.w * (uint32_t)sizeof(T), &bar);
    };
    // one thread: ask L2 for the rows of a coming item (its twin rows are read by plain loads, the upright rows by the bulk copy)
    auto warm_l2 = [&](const int4 &it) {
        if (!has_pairs(it)) return;
        const int r0 = it.w & 0xffff, r1 = it.w >> 16;
        const uint32_t bytes = (uint32_t)(r1 - r0) * a.w * (uint32_t)sizeof(T);
        bulk_prefetch_l2(twin_of(it.x, item_limb(it)) + (size_t)r0 * a.w, bytes);
        bulk_prefetch_l2(upright_of(it.x, item_limb(it)) + (size_t)r0 * a.w, bytes);
    };

    pdl_wait();  // launched as a programmatic dependent: the work queue and the band lists come from the kernels before
    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
        const int4 first = a.item_meta[blockIdx.x];  // the grid never exceeds the item count
        upright_rows(first);
        s_next = first;
    }
    cubic_tab4_init(tab4);
    __syncthreads();

    uint32_t up_uses = 0;  // completed waits on the barrier (one per item with pairs)
    int bset = 0;          // the set of band counts in use
    int pre_img = -1;      // image whose counts are on their way into the other set (asynchronous copies), or -1
    int pending_img = -1;  // image of the limb this CTA finished last and has not published yet (see limbs_stream_kernel)
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
        const int r0 = cur.w & 0xffff, r1 = cur.w >> 16;
        unsigned claimed = 0;
        int4 nmeta = make_int4(-1, 0, 0, 0);
        // thread 0: who is next (one load, issued early), and — once the plane is dead — start its upright plane
        auto resolve_next = [&]() {
            if (tid == 0) {
                const int nitem = claim_to_item(claimed);
                if (nitem >= 0) {
                    nmeta = a.item_meta[nitem];
                    warm_l2(nmeta);
                }
                s_nimg = (nitem >= 0 && item_limb(nmeta) < kLimbs) ? nmeta.x : -1;
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
            // ---- group the image (pafprocess.cpp:133-285) once its 30 limbs are in; no plane copy is in flight: the tables
            //      take the whole allocation ----
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
                group_image(a.group, img, smem_raw, (size_t)a.smem_bytes, a.timeline ? a.timeline + ((size_t)img * 31 + 30) * 4 + 2 : nullptr);
                limb_stamp(a, img, 30, 1);
            }
            if (warp == 0) __syncwarp();  // the chain ran on warp 0: its lanes are done with the tables before a copy may land there
            // claim only now: nothing must wait behind a grouping chain
            if (tid == 0) claimed = atomicAdd(a.work_counter, 1u);
            resolve_next();
            prefetch_next();
            __syncthreads();
            continue;
        }
        const int pa = c_limb_a[limb], pb = c_limb_b[limb];
        const bool pairs = nA > 0 && nB > 0;
        const T *twin = twin_of(img, limb);
        if (tid == 0) claimed = atomicAdd(a.work_counter, 1u);  // next claim out now, answer used later
        if (tid < kL3Pairs / 32) s_accept[tid] = 0;
        limb_stamp(a, img, limb, 0);
        const int nbitems = kParts * a.nbands;
        if (pre_img == img) {  // band counts / first ids of this image: copied under the item before
            bset ^= 1;
            asm volatile("cp.async.wait_all;" ::: "memory");
        } else {
            for (int i = tid; i < nbitems; i += kLimbThreads) s_bcnt[bset][i] = a.band_count[(size_t)img * nbitems + i];
            if (tid <= kParts) s_pbase[bset][tid] = a.partbase[img * (kParts + 1) + tid];
        }
        pre_img = -1;
        const int *s_bc = s_bcnt[bset], *s_partbase = s_pbase[bset];
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
            // ---- the twin rows: the first batch of loads is out before anything else of this item waits ----
            const int wv = a.w / V, total = (r1 - r0) * wv;
            const int shift = ((wv & (wv - 1)) == 0) ? (__ffs(wv) - 1) : -1;
            const Vec *t4 = reinterpret_cast<const Vec *>(twin + (size_t)r0 * a.w);
            auto twin_piece = [&](int i) {  // the piece that mirrors piece i of the rows: same row, column wv - 1 - col
                if (shift >= 0) return __ldcs(t4 + (i ^ (wv - 1)));
                const int r = i / wv, col = i - r * wv;
                return __ldcs(t4 + r * wv + (wv - 1 - col));
            };
            Vec tcur[kL3FoldU], tnxt[kL3FoldU];
#pragma unroll
            for (int k = 0; k < kL3FoldU; k++) {
                const int i = tid + k * kLimbThreads;
                if (i < total) tcur[k] = twin_piece(i);
            }
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
            // ---- fold the twin rows into the upright plane (utils/parse_skeletons.py:91-92) ----
            mbar_wait(&bar, up_uses & 1u);
            up_uses++;
            {
                Vec *o4 = reinterpret_cast<Vec *>(plane_s) + (size_t)r0 * wv;
                for (int base = tid; base < total; base += kL3FoldU * kLimbThreads) {
#pragma unroll
                    for (int k = 0; k < kL3FoldU; k++) {
                        const int i = base + (kL3FoldU + k) * kLimbThreads;
                        if (i < total) tnxt[k] = twin_piece(i);
                    }
#pragma unroll
                    for (int k = 0; k < kL3FoldU; k++) {
                        const int i = base + k * kLimbThreads;
                        if (i < total) {
                            Vec v = o4[i];
                            fold_mirrored(v, tcur[k]);
                            o4[i] = v;
                        }
                    }
#pragma unroll
                    for (int k = 0; k < kL3FoldU; k++) tcur[k] = tnxt[k];
                }
            }
            __syncthreads();  // the plane and the lists are complete
            limb_stamp(a, img, limb, 1);
            {  // the next item's band counts, copied asynchronously under the scoring
                const int nimg = s_nimg;
                if (nimg >= 0) {
                    for (int i = tid; i < nbitems; i += kLimbThreads) cp_async4(&s_bcnt[bset ^ 1][i], a.band_count + (size_t)nimg * nbitems + i);
                    if (tid <= kParts) cp_async4(&s_pbase[bset ^ 1][tid], a.partbase + nimg * (kParts + 1) + tid);
                }
                pre_img = nimg;
            }
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
            const int npairs = (a.debug & 2) ? 0 : nA * nB;
            if (npairs > kL3Pairs) {
                // a crowd: one thread per pair does its own samples (score_and_link's direct path), candidates in global memory
                m.keys = keys_g;
                m.cand_limb = cand_limb_g;
                m.cand_pair = cand_pair_g;
                n_links = score_and_link<kLimbThreads>(a, m, smp, img, limb, nA, nB, prefetch_next);
            } else {
                m.keys = keys_s;
                m.cand_limb = cand_limb_s;
                m.cand_pair = cand_pair_s;
                // ---- lane q of warp w owns pair 8q + w: geometry (pafprocess.cpp:66-73) ----
                const int p_mine = lane * (kLimbThreads / 32) + warp;
                const bool mine = p_mine < npairs;
                int ia = 0, ib = 0;
                DevPeak A, B;
                PairGeo g;
                g.steps = 0;
                g.ax = g.ay = g.sx = g.sy = 0.f;
                if (mine) {
                    ia = p_mine / nB;
                    ib = p_mine - ia * nB;
                    A = m.sA[ia];
                    B = m.sB[ib];
                    g = pair_geometry(A, B, a.max_steps);
                }
                // ---- the warp walks its pairs in groups of `gsz`: the (pair, sample) points of a group are spread over the
                //      lanes (:76-84), the values go through the warp's own 512 bytes of shared memory (the candidate arrays, idle
                //      until the compaction below), then lane j adds the samples of the group's j-th pair in order (:85-88) ----
                float my_sum = 0.0f;
                int my_above = 0;
                const int nslots = npairs > warp ? (npairs - warp + kLimbThreads / 32 - 1) / (kLimbThreads / 32) : 0;
                float *wbuf = reinterpret_cast<float *>(keys_s) + warp * kL3WarpVals;
                const int ms = a.max_steps, gsz = min(32, kL3WarpVals / ms);
                const unsigned rcp = 65536u / (unsigned)ms + 1u;  // f / ms == (f * rcp) >> 16 for f < 256
                for (int q0 = 0; q0 < nslots; q0 += gsz) {
                    const int gp = min(gsz, nslots - q0), nsamp = gp * ms;
                    for (int fb = 0; fb < nsamp; fb += 32) {
                        const int f = fb + lane;
                        const int j = (int)(((unsigned)f * rcp) >> 16), i = f - j * ms;
                        const int src = min(q0 + j, 31);
                        PairGeo gq;
                        gq.ax = __shfl_sync(0xffffffffu, g.ax, src);
                        gq.ay = __shfl_sync(0xffffffffu, g.ay, src);
                        gq.sx = __shfl_sync(0xffffffffu, g.sx, src);
                        gq.sy = __shfl_sync(0xffffffffu, g.sy, src);
                        gq.steps = __shfl_sync(0xffffffffu, g.steps, src);
                        if (f < nsamp && i < gq.steps) {
                            int px, py;
                            pair_sample_xy(gq, i, px, py);
                            wbuf[f] = smp.at(px, py);
                        }
                    }
                    __syncwarp();
                    const int st = __shfl_sync(0xffffffffu, g.steps, min(q0 + lane, 31));
                    float sum = 0.0f;
                    int above = 0;
                    if (lane < gp) {
                        const float *v = wbuf + lane * ms;
                        for (int i = 0; i < st; i++) {
                            const float vi = v[i];
                            sum = __fadd_rn(sum, vi);  // :85
                            above += (vi > a.thr);     // :87
                        }
                    }
                    const float t_sum = __shfl_sync(0xffffffffu, sum, (lane - q0) & 31);
                    const int t_above = __shfl_sync(0xffffffffu, above, (lane - q0) & 31);
                    if (lane >= q0 && lane < q0 + gp) {
                        my_sum = t_sum;
                        my_above = t_above;
                    }
                    __syncwarp();  // the next group overwrites the values
                }
                // ---- criteria of the lane's own pair (:90-96) ----
                bool ok = false;
                float ls = 0.f, rk = 0.f;
                if (mine && g.steps > 0) {
                    const float len = limb_length(A, B);
                    // :90  float + min(0.0, 0.5*min_img_size/len - 1.0) evaluated in double, narrowed on assignment
                    const double pen_raw = __dsub_rn(__ddiv_rn(__dmul_rn(0.5, (double)a.min_img_size), (double)len), 1.0);
                    const double pen = (pen_raw < 0.0) ? pen_raw : 0.0;
                    ls = __double2float_rn(__dadd_rn((double)__fdiv_rn(my_sum, __int2float_rn(g.steps)), pen));
                    const float need = __fmul_rn(__int2float_rn(g.steps), a.step_ratio);  // :91
                    if (__int2float_rn(my_above) > need && ls > 0.0f) {                   // :93
                        ok = true;
                        rk = __fadd_rn(__fadd_rn(__fmul_rn(a.w_limb, ls), __fmul_rn(a.w_a, A.score)), __fmul_rn(a.w_b, B.score));  // :94-96
                        atomicOr(&s_accept[p_mine >> 5], 1u << (p_mine & 31));
                    }
                }
                __syncthreads();  // every sample has been read: the plane is dead
                prefetch_next();
                limb_stamp(a, img, limb, 2);
                // ---- a candidate's place in pair order = accepted pairs below its own ----
                int ncand = 0, idx = 0;
#pragma unroll
                for (int wd = 0; wd < kL3Pairs / 32; wd++) {
                    const unsigned bits = s_accept[wd];
                    ncand += __popc(bits);
                    if (wd < (p_mine >> 5)) idx += __popc(bits);
                    else if (wd == (p_mine >> 5)) idx += __popc(bits & ((1u << (p_mine & 31)) - 1u));
                }
                if (ok && idx < a.cand_cap) {
                    keys_s[idx] = ((uint64_t)float_orderable(rk) << 32) | (uint32_t)idx;
                    cand_limb_s[idx] = ls;
                    cand_pair_s[idx] = ((uint32_t)ia << 16) | (uint32_t)ib;
                }
                __syncthreads();
                n_links = order_and_link<kLimbThreads, kL3Pairs>(a, m, img, limb, nA, nB, ncand);
            }
        }
        if (tid == 0) a.link_count[img * kLimbs + limb] = n_links;
        limb_stamp(a, img, limb, 3);

        // ---- connections and count of this limb are out (warp 0 wrote them); the image's group item counts 30 tickets ----
        pending_img = img;
        __syncthreads();  // s_next, and every use of this round's shared data
    }
    publish_pending();
}

}  // namespace ibp
