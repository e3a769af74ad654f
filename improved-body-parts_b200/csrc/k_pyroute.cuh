// The reference's PYTHON assignment route in float64: find_connections (utils/parse_skeletons.py:324-410) and
// find_humans (utils/parse_skeletons.py:413-599) — what evaluate.py runs without --run_cpp (evaluate.py:88-90), the
// only grouping the original route (predict + find_peaks) has.  It differs from the C++ route in a dozen places
// (SURVEY.md §8a "C++ vs Python"); the ones that shape the code below:
//   * peak positions are float64 sub-pixel values; sample points come from np.linspace and np.round (half to even);
//   * every score is float64; the mean of the samples is numpy's pairwise sum / n;
//   * the candidate order is Python's stable sort: overall score descending, ties in pair order;
//   * a third skeleton matching a connection is ignored instead of dropping the connection;
//   * "has this part" is id >= 0; merging takes the element-wise maximum; merge needs score AND length to agree;
//   * the replace branch subtracts the OLD joint's contribution before adding the new one;
//   * peak scores are looked up as joint_candidates[int(peak_id)]: the concatenated list indexed by the id VALUE.
// Limb values are read from the image-resolution float64 planes of predict() by nearest-pixel gather (L2-resident
// per image: 30 x H x W x 8 bytes), one CTA per (image, limb) for connections, one warp per image for people.
#pragma once
#include "ibp_device.cuh"

namespace ibp {

struct PyParams {  // utils/config:17-27 through config_reader.py, as the Python floats they are
    double thre2;            // 0.1
    double connect_ration;   // 0.8
    int mid_num;             // 20
    int len_rate;            // 16
    double connection_tole;  // 0.7
    double min_score;        // 0.45 literal at parse_skeletons.py:594
    int min_parts;           // 2 literal at parse_skeletons.py:594
};

struct PyConnArgs {
    const DevPeak *peaks;     // (batch, 18, cap): id and float32 score
    const double *subpix;     // (batch, 18, cap, 2): float64 x, y
    const double *score64;    // (batch, 18, cap) float64 scores, or null: use (double)peaks.score
    const int *part_count;    // (batch, 18)
    int cap;
    const double *paf;        // (batch, paf_chan_stride, H, W) planar float64, limb k in channel k
    int paf_chan_stride;
    int H, W;
    int img_height;
    PyParams P;
    int cand_cap;
    double *conn;             // (batch, 30, cap, 6) rows [src_id, dst_id, score, i, j, length]
    int *conn_count;          // (batch, 30)
    double *cand_score;       // (batch, line_stride) scores in concatenation order, written by the limb-0 CTA, or null
    int line_stride;
    uint32_t *status;         // (batch)
};

constexpr int kPyThreads = 256;
constexpr int kPyMaxSteps = 32;  // mid_num <= 32

struct PyPeak {
    double x, y, score;
    int id, pad;
};

// numpy pairwise sum (loops_utils.h) for n <= 32 float64 values held in a local array
__device__ __forceinline__ double py_pairwise_sum(const double *a, int n) {
    if (n < 8) {
        double res = 0.;
        for (int i = 0; i < n; i++) res = __dadd_rn(res, a[i]);
        return res;
    }
    double r[8];
#pragma unroll
    for (int k = 0; k < 8; k++) r[k] = a[k];
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
#pragma unroll
        for (int k = 0; k < 8; k++) r[k] = __dadd_rn(r[k], a[i + k]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])), __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; i++) res = __dadd_rn(res, a[i]);
    return res;
}

// np.linspace(start, stop, num)[i] (numpy/_core/function_base.py): i*step + start with step = (stop-start)/(num-1),
// (i/div)*delta + start when step == 0, the last element forced to `stop`, a single element = start
__device__ __forceinline__ double py_linspace_at(double start, double stop, int num, int i) {
    const int div = num - 1;
    const double delta = __dsub_rn(stop, start);
    if (div <= 0) return __dadd_rn(__dmul_rn((double)i, delta), start);
    if (i == num - 1) return stop;
    const double step = __ddiv_rn(delta, (double)div);
    if (step == 0.0) return __dadd_rn(__dmul_rn(__ddiv_rn((double)i, (double)div), delta), start);
    return __dadd_rn(__dmul_rn((double)i, step), start);
}

// ---------------------------------------------------------------------------------------------------------------
// find_connections: one CTA per (image, limb); one thread per candidate pair, compaction in pair order, stable rank
// sort, greedy pick by warp 0.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPyThreads) py_connections_kernel(PyConnArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int limb = blockIdx.x, img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int pa = c_limb_a[limb], pb = c_limb_b[limb];
    const int *pc = a.part_count + img * kParts;
    const int nA = pc[pa], nB = pc[pb];

    PyPeak *sA = reinterpret_cast<PyPeak *>(smem_raw);
    PyPeak *sB = sA + a.cap;
    double *c_overall = reinterpret_cast<double *>(sB + a.cap);  // [cand_cap]
    double *c_score = c_overall + a.cand_cap;
    double *c_len = c_score + a.cand_cap;
    uint32_t *c_pair = reinterpret_cast<uint32_t *>(c_len + a.cand_cap);
    int *c_order = reinterpret_cast<int *>(c_pair + a.cand_cap);
    unsigned char *used_a = reinterpret_cast<unsigned char *>(c_order + a.cand_cap);
    unsigned char *used_b = used_a + a.cap;
    __shared__ int wcnt[kPyThreads / 32];

    auto load_list = [&](PyPeak *dst, unsigned char *used, int part, int n) {
        const size_t base = ((size_t)img * kParts + part) * a.cap;
        for (int i = tid; i < n; i += kPyThreads) {
            PyPeak p;
            p.x = a.subpix[(base + i) * 2];
            p.y = a.subpix[(base + i) * 2 + 1];
            p.score = a.score64 ? a.score64[base + i] : (double)a.peaks[base + i].score;
            p.id = a.peaks[base + i].id;
            p.pad = 0;
            dst[i] = p;
            used[i] = 0;
        }
    };
    load_list(sA, used_a, pa, nA);
    load_list(sB, used_b, pb, nB);
    // joint_candidates[:, 2] in concatenation order (parse_skeletons.py:427), for find_humans' lookups
    if (limb == 0 && a.cand_score) {
        int pos0 = 0;
        for (int p = 0; p < kParts; p++) {
            const int n = pc[p];
            const size_t base = ((size_t)img * kParts + p) * a.cap;
            for (int i = tid; i < n; i += kPyThreads)
                if (pos0 + i < a.line_stride)
                    a.cand_score[(size_t)img * a.line_stride + pos0 + i] = a.score64 ? a.score64[base + i] : (double)a.peaks[base + i].score;
            pos0 += n;
        }
    }
    __syncthreads();

    const double *plane = a.paf + ((size_t)img * a.paf_chan_stride + limb) * (size_t)a.H * a.W;
    const int npairs = nA * nB;
    int ncand = 0;
    for (int base = 0; base < npairs; base += kPyThreads) {
        const int p = base + tid;
        bool ok = false;
        double score = 0., length = 0., overall = 0.;
        int ia = 0, ib = 0;
        if (p < npairs) {
            ia = p / nB;
            ib = p - ia * nB;
            const PyPeak A = sA[ia], B = sB[ib];
            const double dx = __dsub_rn(B.x, A.x), dy = __dsub_rn(B.y, A.y);
            length = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));           // l.353
            if (length != 0.0) {                                                          // l.357
                const int n = min((int)__double2ll_rn(__dadd_rn(length, 1.0)), a.P.mid_num);  // l.354 round(): half to even
                double vals[kPyMaxSteps];
                int above = 0;
                for (int i = 0; i < n; i++) {
                    // np.round = rint (half to even); the reference would wrap / fault outside the map: clamp instead
                    const int xi = clampi((int)__double2ll_rn(py_linspace_at(A.x, B.x, n, i)), 0, a.W - 1);   // l.361
                    const int yi = clampi((int)__double2ll_rn(py_linspace_at(A.y, B.y, n, i)), 0, a.H - 1);   // l.362
                    const double v = __ldg(plane + (size_t)yi * a.W + xi);
                    vals[i] = v;
                    above += (v > a.P.thre2);
                }
                const double mean = __ddiv_rn(py_pairwise_sum(vals, n), (double)n);
                const double pen_raw = __dsub_rn(__ddiv_rn(__dmul_rn(0.5, (double)a.img_height), length), 1.0);
                score = __dadd_rn(mean, pen_raw < 0.0 ? pen_raw : 0.0);                      // l.367
                if ((double)above > __dmul_rn((double)n, a.P.connect_ration) && score > 0.0) {   // l.373-378
                    ok = true;
                    overall = __dadd_rn(__dadd_rn(__dmul_rn(0.5, score), __dmul_rn(0.25, A.score)), __dmul_rn(0.25, B.score));  // l.381
                }
            }
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) wcnt[warp] = __popc(bal);
        __syncthreads();
        int woff = 0, tot = 0;
#pragma unroll
        for (int i = 0; i < kPyThreads / 32; i++) {
            const int t = wcnt[i];
            if (i < warp) woff += t;
            tot += t;
        }
        if (ok) {
            const int idx = ncand + woff + __popc(bal & ((1u << lane) - 1u));
            if (idx < a.cand_cap) {
                c_overall[idx] = overall;
                c_score[idx] = score;
                c_len[idx] = length;
                c_pair[idx] = ((uint32_t)ia << 16) | (uint32_t)ib;
            }
        }
        ncand += tot;
        __syncthreads();
    }
    if (ncand > a.cand_cap) {
        if (tid == 0) atomicOr(a.status + img, IBP_ST_CAND_OVERFLOW);
        ncand = a.cand_cap;
    }
    // stable descending order (l.391: sorted(..., reverse=True) keeps equal keys in input order)
    for (int t = tid; t < ncand; t += kPyThreads) {
        const double mine = c_overall[t];
        int rank = 0;
        for (int j = 0; j < ncand; j++) {
            const double o = c_overall[j];
            rank += (o > mine) || (o == mine && j < t);
        }
        c_order[rank] = t;
    }
    __syncthreads();
    // greedy pick (l.393-407)
    if (warp == 0) {
        const int limit = min(nA, nB);
        double *out = a.conn + ((size_t)img * kLimbs + limb) * a.cap * 6;
        int nl = 0;
        for (int base = 0; base < ncand && nl < limit; base += 32) {
            const int i = base + lane;
            bool valid = i < ncand;
            int ia = 0, ib = 0, cidx = 0;
            if (valid) {
                cidx = c_order[i];
                const uint32_t pr = c_pair[cidx];
                ia = (int)(pr >> 16);
                ib = (int)(pr & 0xffffu);
            }
            while (nl < limit) {
                const bool free_now = valid && !used_a[ia] && !used_b[ib];
                const uint32_t bal = __ballot_sync(0xffffffffu, free_now);
                if (bal == 0) break;
                const int leader = __ffs(bal) - 1;
                if (lane == leader) {
                    used_a[ia] = 1;
                    used_b[ib] = 1;
                    double *row = out + (size_t)nl * 6;
                    row[0] = (double)sA[ia].id;
                    row[1] = (double)sB[ib].id;
                    row[2] = c_score[cidx];
                    row[3] = (double)ia;
                    row[4] = (double)ib;
                    row[5] = c_len[cidx];
                    valid = false;
                }
                nl++;
                __syncwarp();
            }
        }
        if (lane == 0) a.conn_count[img * kLimbs + limb] = nl;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// find_humans: one warp per image.  Table row = (20,2) float64: [part] = (peak id, limb score), [18] = (total, -1),
// [19] = (count, last length).  np.delete keeps the order of the remaining rows, so a merged-away row is only marked
// (ids that match nothing, count -1) and skipped by the final filter.
// ---------------------------------------------------------------------------------------------------------------
struct PyHumanArgs {
    const double *conn;        // (batch, 30, cap, 6)
    const int *conn_count;     // (batch, 30)
    const int *part_count;     // (batch, 18): a limb whose two lists are both empty is "special" (l.339-341) — it has
                               // no rows either way, so the count alone decides
    int cap;
    const double *cand_score;  // (batch, line_stride) joint_candidates[:, 2]
    int line_stride;
    PyParams P;
    int max_bodies, max_humans;
    double *people;            // (batch, max_humans, 20, 2)
    int *n_people;             // (batch)
    uint32_t *status;          // (batch)
};

__global__ void __launch_bounds__(32) py_humans_kernel(PyHumanArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int img = blockIdx.x, lane = threadIdx.x;
    double *tab = reinterpret_cast<double *>(smem_raw);  // [S][20][2]
    const int S = a.max_bodies;
    const double *cand = a.cand_score + (size_t)img * a.line_stride;
    auto score_of = [&](double id) { return cand[clampi((int)id, 0, a.line_stride - 1)]; };  // joint_candidates[int(id), 2]
#define ROW(s, k, c) tab[((s)*20 + (k)) * 2 + (c)]
    int nb = 0;
    bool overflow = false;
    const double rate = (double)a.P.len_rate;
    for (int limb = 0; limb < kLimbs; limb++) {
        const int pa = c_limb_a[limb], pb = c_limb_b[limb];
        const int nl = a.conn_count[img * kLimbs + limb];
        const double *rows = a.conn + ((size_t)img * kLimbs + limb) * a.cap * 6;
        for (int li = 0; li < nl; li++) {
            const double src_id = rows[li * 6 + 0], dst_id = rows[li * 6 + 1], score = rows[li * 6 + 2], length = rows[li * 6 + 5];
            // l.441-449: the first two matching rows; further matches are ignored
            int hits = 0, first = 0, second = 0;
            for (int base = 0; base < nb; base += 32) {
                const int s = base + lane;
                const bool m = (s < nb) && (ROW(s, pa, 0) == src_id || ROW(s, pb, 0) == dst_id);
                uint32_t bal = __ballot_sync(0xffffffffu, m);
                while (bal && hits < 2) {
                    const int pos = __ffs(bal) - 1;
                    bal &= bal - 1;
                    if (hits == 0) first = base + pos; else second = base + pos;
                    hits++;
                }
            }
            if (hits == 1) {
                if (lane == 0) {
                    const int f = first;
                    const double cur_id = ROW(f, pb, 0), cur_sc = ROW(f, pb, 1), reach = ROW(f, 19, 1);
                    const bool len_ok = __dmul_rn(reach, rate) > length;
                    if ((int)cur_id == -1 && len_ok) {                                                // l.458-471
                        ROW(f, pb, 0) = dst_id;
                        ROW(f, pb, 1) = score;
                        ROW(f, 19, 0) = __dadd_rn(ROW(f, 19, 0), 1.0);
                        ROW(f, 19, 1) = length > reach ? length : reach;
                        ROW(f, 18, 0) = __dadd_rn(ROW(f, 18, 0), __dadd_rn(score_of(dst_id), score));
                    } else if (((int)cur_id != (int)dst_id && cur_sc <= score && len_ok) ||
                               ((int)cur_id == (int)dst_id && cur_sc <= score)) {                      // l.473-489
                        const double tot = __dsub_rn(ROW(f, 18, 0), __dadd_rn(score_of(cur_id), cur_sc));
                        ROW(f, pb, 0) = dst_id;
                        ROW(f, pb, 1) = score;
                        ROW(f, 19, 1) = length > reach ? length : reach;
                        ROW(f, 18, 0) = __dadd_rn(tot, __dadd_rn(score_of(dst_id), score));
                    }
                }
                __syncwarp();
            } else if (hits == 2) {  // l.491-524
                const int f = first, g = second;
                double id1 = -1., id2 = -1., sc1 = 0., sc2 = 0.;
                if (lane < kParts) {
                    id1 = ROW(f, lane, 0);
                    sc1 = ROW(f, lane, 1);
                    id2 = ROW(g, lane, 0);
                    sc2 = ROW(g, lane, 1);
                }
                const bool m1 = lane < kParts && id1 >= 0.0, m2 = lane < kParts && id2 >= 0.0;
                if (__ballot_sync(0xffffffffu, m1 && m2) == 0) {
                    // np.min over the member scores of each person, then the smaller of the two (l.508-510)
                    double lo = (m1 || m2) ? ((m1 && m2) ? fmin(sc1, sc2) : (m1 ? sc1 : sc2)) : INFINITY;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                    const double reach = ROW(f, 19, 1);
                    if (score >= __dmul_rn(a.P.connection_tole, lo) && length < __dmul_rn(reach, rate)) {   // l.511-512
                        if (lane < kParts) {
                            ROW(f, lane, 0) = id1 > id2 ? id1 : id2;  // np.maximum (l.516)
                            ROW(f, lane, 1) = sc1 > sc2 ? sc1 : sc2;
                            ROW(g, lane, 0) = -3.0;                    // np.delete (l.524): matches nothing from now on
                        }
                        if (lane == 0) {
                            ROW(f, 19, 0) = __dadd_rn(ROW(f, 19, 0), ROW(g, 19, 0));
                            ROW(f, 19, 1) = length > reach ? length : reach;
                            ROW(f, 18, 0) = __dadd_rn(ROW(f, 18, 0), __dadd_rn(ROW(g, 18, 0), score));
                            ROW(g, 19, 0) = -1.0;
                        }
                    }
                }
                // shared joints: remove_recon = 0 (utils/config:27) => nothing (l.526-566)
                __syncwarp();
            } else {  // l.570-585: a new person
                if (nb < S) {
                    if (lane < 20) {
                        const bool end = (lane == pa || lane == pb);
                        ROW(nb, lane, 0) = lane == pa ? src_id : (lane == pb ? dst_id : -1.0);
                        ROW(nb, lane, 1) = end ? score : -1.0;
                    }
                    __syncwarp();
                    if (lane == 0) {
                        ROW(nb, 19, 0) = 2.0;
                        ROW(nb, 19, 1) = length;
                        // sum(joint_candidates[ids, 2]) + score: Python's sum starts from int 0
                        ROW(nb, 18, 0) = __dadd_rn(__dadd_rn(score_of(src_id), score_of(dst_id)), score);
                    }
                    nb++;
                } else {
                    overflow = true;
                }
                __syncwarp();
            }
        }
    }
    // l.591-596 filter, order preserved
    int n_out = 0;
    double *out = a.people + (size_t)img * a.max_humans * 40;
    for (int s = 0; s < nb; s++) {
        const double cnt = ROW(s, 19, 0), tot = ROW(s, 18, 0);
        const bool drop = cnt < (double)a.P.min_parts || __ddiv_rn(tot, cnt) < a.P.min_score;
        if (!drop) {
            if (n_out < a.max_humans)
                for (int i = lane; i < 40; i += 32) out[(size_t)n_out * 40 + i] = tab[s * 40 + i];
            n_out++;
        }
    }
    if (lane == 0) {
        a.n_people[img] = min(n_out, a.max_humans);
        uint32_t st = 0;
        if (overflow) st |= IBP_ST_BODY_OVERFLOW;
        if (n_out > a.max_humans) st |= IBP_ST_HUMAN_OVERFLOW;
        if (st) atomicOr(a.status + img, st);
    }
#undef ROW
}

}  // namespace ibp
