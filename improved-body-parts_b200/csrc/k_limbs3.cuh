// K3 + K4, the streaming limb kernel with THREE CTAs per SM for float32 planes (device-resident network output).
//
// Same work queue, same arithmetic, same phases and the same publish / grouping protocol as limbs_stream_kernel
// (k_limbs.cuh).  What an item costs there is the length of a mostly serial chain (claim -> lists -> plane -> fold ->
// score -> order -> publish): four warps run it about as fast as eight, and the kernel's rate is items in flight x 1 /
// chain length.  limbs_stream_kernel keeps two items per SM because each carries 36 KB of scratch next to its 64 KB
// plane.  Here an item holds 64 KB + 9.6 KB, so three 128-thread CTAs fit an SM:
//   * the ring the twin rows stream through lies in the rows of the plane buffer the item does NOT use — only the
//     feature rows [r0, r1) a sample can touch are staged, the larger of [0, r0) and [r1, h) takes the ring (an item that
//     needs nearly every row falls back to a small ring in the candidate / sample scratch);
//   * the two-phase scoring takes 24 pairs per round instead of 128 (2.4 KB of geometry and sample values);
//   * 128 candidates in shared memory — the two-phase path has at most 128 pairs; a crowded limb (one thread per pair) keeps
//     its candidates in a per-CTA scratch in global memory, and so does the introsort emulation;
//   * the grouping tables use the plane's space (no copy is in flight while an image is grouped).
#pragma once
#include "k_limbs.cuh"

namespace ibp {

constexpr int kL3Threads = 128;
constexpr int kL3Chunk = 24;  // pairs per round of the two-phase scoring

// shared memory of a CTA beyond the plane: two peak lists with their flags and scores, 128 candidates, one round of pairs
__host__ __device__ inline size_t l3_scratch_bytes(int cap, int max_steps) {
    size_t off = (size_t)cap * (2 * sizeof(DevPeak) + 2);
    off = (off + 15) & ~(size_t)15;
    off += (size_t)cap * 8;
    off = (off + 15) & ~(size_t)15;
    off += (size_t)kL3Threads * 16;
    off += (size_t)kL3Chunk * (sizeof(PairGeo) + (size_t)max_steps * 4);
    return (off + 15) & ~(size_t)15;
}
// ... of which the ring of last resort (candidates + round scratch)
__host__ __device__ inline size_t l3_fallback_ring_bytes(int max_steps) {
    return ((size_t)kL3Threads * 16 + (size_t)kL3Chunk * (sizeof(PairGeo) + (size_t)max_steps * 4)) & ~(size_t)15;
}
// global scratch of a CTA: keys, limb scores, pairs of cand_cap candidates + the emulation array
__host__ __device__ inline size_t l3_global_scratch_bytes(int cand_cap) { return (size_t)cand_cap * 24; }

template <typename T>
__global__ void __launch_bounds__(kL3Threads, 3) limbs_stream3_kernel(LimbArgs a) {
    constexpr int kLimbThreads = kL3Threads;
    constexpr int V = 16 / (int)sizeof(T);  // elements per 16-byte vector
    using Vec = typename Vec16<T>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int items = a.batch * (kLimbs + 1);
    const size_t hw = (size_t)a.h * a.w;

    // ---- carve shared memory: plane | peak lists | candidates | one round of pairs ----
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
    const size_t fb_off = off;  // the ring of last resort starts here
    uint64_t *keys_s = reinterpret_cast<uint64_t *>(smem_raw + off);
    off += (size_t)kLimbThreads * 8;
    float *cand_limb_s = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)kLimbThreads * 4;
    uint32_t *cand_pair_s = reinterpret_cast<uint32_t *>(smem_raw + off);
    off += (size_t)kLimbThreads * 4;
    m.geo = reinterpret_cast<PairGeo *>(smem_raw + off);
    off += (size_t)kL3Chunk * sizeof(PairGeo);
    m.vals = reinterpret_cast<float *>(smem_raw + off);
    off += (size_t)kL3Chunk * a.max_steps * 4;
    T *fb_ring = reinterpret_cast<T *>(smem_raw + fb_off);
    const int row_bytes = a.w * (int)sizeof(T);
    const int fb_slot_rows = (int)(((off - fb_off) & ~(size_t)15) / kLimbRing) / row_bytes;  // >= 1 (the host checks)
    unsigned char *gs = a.cta_scratch + (size_t)blockIdx.x * a.cta_scratch_bytes;
    uint64_t *keys_g = reinterpret_cast<uint64_t *>(gs);
    float *cand_limb_g = reinterpret_cast<float *>(gs + (size_t)a.cand_cap * 8);
    uint32_t *cand_pair_g = reinterpret_cast<uint32_t *>(gs + (size_t)a.cand_cap * 12);
    m.emu = reinterpret_cast<uint64_t *>(gs + (size_t)a.cand_cap * 16);
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
    auto item_limb = [](const int4 &it) { return it.z & 0xff; };
    auto has_pairs = [&](const int4 &it) { return (it.y & 0xffff) > 0 && (it.y >> 16) > 0 && item_limb(it) < kLimbs; };
    // one thread: the feature rows [r0, r1) of a limb item's upright plane straight to their place (item.w = r0 | r1 << 16)
    auto upright_rows = [&](const int4 &it) {
        if (!has_pairs(it)) return;
        const int r0 = it.w & 0xffff, r1 = it.w >> 16;
        bulk_rows(plane_s + (size_t)r0 * a.w, upright_of(it.x, item_limb(it)) + (size_t)r0 * a.w, (uint32_t)(r1 - r0) * (uint32_t)row_bytes, &bars[0]);
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
    int pending_img = -1;   // image of the limb this CTA finished last and has not published yet (see limbs_stream_kernel)
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
            if (tid < 32) __syncwarp();  // the chain ran on warp 0: its lanes are done with the tables before a copy may land there
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
        // the ring of this item: the larger run of plane rows it does not use, or the scratch if that is too small
        const int free_rows = max(r0, a.h - r1);
        const bool in_plane = free_rows / kLimbRing >= fb_slot_rows;
        T *ring = in_plane ? plane_s + (size_t)(r0 >= a.h - r1 ? 0 : r1) * a.w : fb_ring;
        const int slot_rows = in_plane ? free_rows / kLimbRing : fb_slot_rows;
        const int nchunks = (r1 - r0 + slot_rows - 1) / slot_rows;
        auto twin_chunk = [&](int c) {  // one thread: chunk c of the twin rows into its ring slot
            const int s = c % kLimbRing, first = r0 + c * slot_rows;
            const int rows = min(slot_rows, r1 - first);
            bulk_rows(ring + (size_t)s * slot_rows * a.w, twin + (size_t)first * a.w, (uint32_t)rows * (uint32_t)row_bytes, &bars[1 + s]);
        };
        // thread 0: next claim out now, answer used later; the ring is free again: first twin chunks out
        if (tid == 0) {
            claimed = atomicAdd(a.work_counter, 1u);
            if (pairs) {
                fence_proxy_async();  // the previous round's generic-proxy use of the ring region comes first
                for (int c = 0; c < kLimbRing && c < nchunks; c++) twin_chunk(c);
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
                const int shift = ((wv & (wv - 1)) == 0) ? (__ffs(wv) - 1) : -1;
                Vec *d4 = reinterpret_cast<Vec *>(plane_s);
                for (int c = 0; c < nchunks; c++) {
                    const int s = c % kLimbRing;
                    mbar_wait(&bars[1 + s], (ring_par >> s) & 1u);
                    ring_par ^= 1u << s;
                    const int first = r0 + c * slot_rows;
                    const int rows = min(slot_rows, r1 - first);
                    const Vec *r4 = reinterpret_cast<const Vec *>(ring + (size_t)s * slot_rows * a.w);
                    Vec *o4 = d4 + (size_t)first * wv;
                    for (int i = tid; i < rows * wv; i += kLimbThreads) {
                        const int mi = (shift >= 0) ? (i ^ (wv - 1)) : ((i / wv) * wv + (wv - 1 - (i - (i / wv) * wv)));
                        Vec v = o4[i];
                        fold_mirrored(v, r4[mi]);
                        o4[i] = v;
                    }
                    if (c + kLimbRing < nchunks) {
                        __syncthreads();  // everybody has read slot s
                        if (tid == 0) {
                            fence_proxy_async();
                            twin_chunk(c + kLimbRing);
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
            if (nA * nB > kLimbThreads) {
                // a crowd: one thread per pair does its own samples (score_and_link's direct path), candidates in global memory
                m.keys = keys_g;
                m.cand_limb = cand_limb_g;
                m.cand_pair = cand_pair_g;
                n_links = score_and_link<kLimbThreads, kL3Chunk, kRankSortMax>(a, m, smp, img, limb, nA, nB, prefetch_next);
            } else {
                m.keys = keys_s;
                m.cand_limb = cand_limb_s;
                m.cand_pair = cand_pair_s;
                n_links = score_and_link<kLimbThreads, kL3Chunk, kLimbThreads>(a, m, smp, img, limb, nA, nB, prefetch_next);
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
