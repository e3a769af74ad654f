// K2 — keypoint peaks: non-maximum suppression, in-order compaction into per-part candidate lists, sub-pixel refine.
//
// Flavour IBP_PEAKS_CROSS_REFINE restates heatmap_nms (utils/parse_skeletons.py:126-176):
//   * find_peaks_refactor (l.106-119): maximum_filter with the 4-connected cross footprint == value, AND value > 0.1
//     (strict).  scipy's default 'reflect' border repeats the edge pixel, so neighbours outside the map never win.
//     np.nonzero order = row-major (y, then x)  ->  peak ids.
//   * per peak (l.141-173): patch [y-2..y+2] x [x-2..x+2] clipped to the map, cv2.resize(patch, fx=4, fy=4, CUBIC)
//     with the border replicated AT THE PATCH EDGE, first arg-max in row-major order; x = 4*x_min + col,
//     y = 4*y_min + row (integer valued), score = the up-sampled maximum.
// Flavour IBP_PEAKS_BOX_CENTROID restates find_peaks (utils/parse_skeletons.py:286-321) with
// util.keypoint_heatmap_nms (utils/util.py:177-185: 3x3 max-pool on a reflect-padded map, value >= thre1, and the
// `heat * keep` product non-zero) and util.refine_centroid (utils/util.py:188-213, float64 moments with numpy's
// summation order); see centroid_refine below.
//
// One CTA per (image, part).  The plane is streamed through shared memory in bands of whole rows (+2 halo rows)
// by the bulk-copy engine — a band of a planar map is one contiguous run of bytes — so HBM sees each value once
// (algorithmic bytes: 18*h*w*4 per image) and the 5-point / 9-point stencil and the 5x5 patches
// are served from shared memory.  Ordered compaction: one ballot word per 32 pixels, a block scan over the
// pop-counts, then every thread emits the set bits of its words — list order equals np.nonzero order.
#pragma once
#include "ibp_device.cuh"

namespace ibp {

struct PeakArgs {
    const float *maps;   // (batch, chan_stride, h, w)
    int chan_stride;
    int first_chan;
    int h, w;
    int rows_per_band;   // owned rows per band
    int rows_buf;        // rows that fit the shared buffer (>= rows_per_band + 2*halo, or h)
    int halo;
    int up;              // up-sampling factor of the refine step
    int win;             // patch half-size / centroid radius
    float thr;
    int cap;             // max peaks per (image, part)
    int flavour;
    DevPeak *peaks;      // (batch, 18, cap)
    float *subpix;       // (batch, 18, cap, 2) float x,y for the centroid flavour (may be null)
    int *part_count;     // (batch, 18)
    uint32_t *status;    // (batch)
};

constexpr int kPeakThreads = 256;

__device__ __forceinline__ int block_exclusive_scan(int v, int *warp_tot, int &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int base = 0, tot = 0;
    const int nw = (blockDim.x + 31) >> 5;
    for (int i = 0; i < nw; i++) {
        const int t = warp_tot[i];
        if (i < warp) base += t;
        tot += t;
    }
    total = tot;
    return base + incl - v;
}

// numpy's pairwise summation for a contiguous run of n < 128 float64 values (numpy/core/src/umath/loops_utils.h
// pairwise_sum): n < 8 -> plain left-to-right loop starting from 0.; otherwise 8 accumulators over blocks of 8,
// combined as ((r0+r1)+(r2+r3)) + ((r4+r5)+(r6+r7)), then the tail added one by one.
__device__ __forceinline__ double np_pairwise_sum_f64(const double *a, int n) {
    if (n < 8) {
        double res = 0.;
        for (int i = 0; i < n; i++) res = __dadd_rn(res, a[i]);
        return res;
    }
    double r[8];
    for (int k = 0; k < 8; k++) r[k] = a[k];
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
        for (int k = 0; k < 8; k++) r[k] = __dadd_rn(r[k], a[i + k]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])), __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; i++) res = __dadd_rn(res, a[i]);
    return res;
}

__global__ void __launch_bounds__(kPeakThreads) peaks_kernel(PeakArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int part = blockIdx.x, img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwarps = kPeakThreads / 32;
    const int w = a.w, h = a.h;

    // carve shared memory
    float *band = reinterpret_cast<float *>(smem_raw);
    size_t off = (size_t)a.rows_buf * w * sizeof(float);
    off = (off + 15) & ~(size_t)15;
    uint32_t *words = reinterpret_cast<uint32_t *>(smem_raw + off);
    const int words_max = (a.rows_per_band * w + 31) / 32;
    off += (size_t)words_max * 4;
    off = (off + 7) & ~(size_t)7;
    const int side = 2 * a.win + 1;
    float *hbuf_all = reinterpret_cast<float *>(smem_raw + off);  // per warp: side * side*up floats
    const int hbuf_per_warp = side * side * a.up;
    off += (size_t)nwarps * hbuf_per_warp * 4;
    off = (off + 7) & ~(size_t)7;
    short2 *pk = reinterpret_cast<short2 *>(smem_raw + off);      // [cap]
    __shared__ uint64_t bar;
    __shared__ int warp_tot[kPeakThreads / 32];

    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    uint32_t phase = 0;

    const float *plane = a.maps + ((size_t)img * a.chan_stride + a.first_chan + part) * (size_t)h * w;
    DevPeak *out = a.peaks + ((size_t)img * kParts + part) * a.cap;
    int n_total = 0;  // peaks found so far (may exceed cap)

    for (int r0 = 0; r0 < h; r0 += a.rows_per_band) {
        const int r1 = min(r0 + a.rows_per_band, h);
        const int l0 = max(r0 - a.halo, 0), l1 = min(r1 + a.halo, h);
        __syncthreads();  // everybody is done with the previous band
        rows_to_smem(band, plane + (size_t)l0 * w, (uint32_t)(l1 - l0) * w, &bar, phase);
#define BAND(y, x) band[((y)-l0) * w + (x)]
        const int npix = (r1 - r0) * w;
        const int npix32 = (npix + 31) & ~31;
        for (int i = tid; i < npix32; i += kPeakThreads) {
            bool keep = false;
            if (i < npix) {
                const int yy = i / w;
                const int x = i - yy * w;
                const int y = r0 + yy;
                const float v = BAND(y, x);
                if (a.flavour == IBP_PEAKS_CROSS_REFINE) {
                    keep = v > a.thr;
                    if (keep) {
                        if (y > 0) keep = keep && (v >= BAND(y - 1, x));
                        if (y < h - 1) keep = keep && (v >= BAND(y + 1, x));
                        if (x > 0) keep = keep && (v >= BAND(y, x - 1));
                        if (x < w - 1) keep = keep && (v >= BAND(y, x + 1));
                    }
                } else {
                    // util.py:181-185: hmax == heat, heat >= thre, and heat*keep != 0
                    keep = (v >= a.thr) && (v != 0.0f);
                    if (keep) {
                        const int ya = max(y - 1, 0), yb = min(y + 1, h - 1), xa = max(x - 1, 0), xb = min(x + 1, w - 1);
                        for (int q = ya; q <= yb; q++)
                            for (int p = xa; p <= xb; p++) keep = keep && (v >= BAND(q, p));
                    }
                }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) words[i >> 5] = bal;
        }
        __syncthreads();
        // scan the pop-counts
        const int nwords = npix32 >> 5;
        const int wpt = (nwords + kPeakThreads - 1) / kPeakThreads;
        const int w_begin = min(tid * wpt, nwords), w_end = min(w_begin + wpt, nwords);
        int cnt = 0;
        for (int q = w_begin; q < w_end; q++) cnt += __popc(words[q]);
        int band_total;
        int pos = n_total + block_exclusive_scan(cnt, warp_tot, band_total);
        for (int q = w_begin; q < w_end; q++) {
            uint32_t bits = words[q];
            while (bits) {
                const int bit = __ffs(bits) - 1;
                bits &= bits - 1;
                const int i = (q << 5) + bit;
                const int yy = i / w;
                if (pos < a.cap) pk[pos] = make_short2((short)(i - yy * w), (short)(r0 + yy));
                pos++;
            }
        }
        __syncthreads();
        const int k_begin = min(n_total, a.cap), k_end = min(n_total + band_total, a.cap);
        n_total += band_total;

        // refine the peaks of this band, one warp per peak
        float *hbuf = hbuf_all + warp * hbuf_per_warp;
        for (int k = k_begin + warp; k < k_end; k += nwarps) {
            const int px = pk[k].x, py = pk[k].y;
            if (a.flavour == IBP_PEAKS_CROSS_REFINE) {
                const int x0 = max(px - a.win, 0), x1 = min(px + a.win, w - 1);
                const int y0 = max(py - a.win, 0), y1 = min(py + a.win, h - 1);
                const int pw = x1 - x0 + 1, ph = y1 - y0 + 1;
                const int W4 = pw * a.up, H4 = ph * a.up;
                // horizontal pass: ph rows x W4 columns
                for (int idx = lane; idx < ph * W4; idx += 32) {
                    const int r = idx / W4, X = idx - r * W4;
                    float xw[4];
                    const int sx = cubic_tap(X, a.up, xw);
                    const int c0 = x0 + clampi(sx - 1, 0, pw - 1), c1 = x0 + clampi(sx, 0, pw - 1);
                    const int c2 = x0 + clampi(sx + 1, 0, pw - 1), c3 = x0 + clampi(sx + 2, 0, pw - 1);
                    const int y = y0 + r;
                    hbuf[r * W4 + X] = hsum4(BAND(y, c0), BAND(y, c1), BAND(y, c2), BAND(y, c3), xw);
                }
                __syncwarp();
                float best = -INFINITY;
                int best_idx = 0x7fffffff;
                for (int idx = lane; idx < H4 * W4; idx += 32) {
                    const int Y = idx / W4, X = idx - Y * W4;
                    float yw[4];
                    const int sy = cubic_tap(Y, a.up, yw);
                    const int q0 = clampi(sy - 1, 0, ph - 1), q1 = clampi(sy, 0, ph - 1);
                    const int q2 = clampi(sy + 1, 0, ph - 1), q3 = clampi(sy + 2, 0, ph - 1);
                    const float v = vsum4(hbuf[q0 * W4 + X], hbuf[q1 * W4 + X], hbuf[q2 * W4 + X], hbuf[q3 * W4 + X], yw);
                    if (v > best) {
                        best = v;
                        best_idx = idx;
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
                    if (ov > best || (ov == best && oi < best_idx)) {
                        best = ov;
                        best_idx = oi;
                    }
                }
                if (lane == 0) {
                    if (best_idx == 0x7fffffff) best_idx = 0;
                    DevPeak p;
                    p.x = x0 * a.up + best_idx % W4;
                    p.y = y0 * a.up + best_idx / W4;
                    p.score = best;
                    p.id = k;
                    out[k] = p;
                }
                __syncwarp();
            } else {
                // util.refine_centroid (util.py:188-213), radius = win
                if (lane == 0) {
                    const int rad = a.win;
                    float fx, fy, sc;
                    if (py + rad + 1 > h || py - rad < 0 || px + rad + 1 > w || px - rad < 0) {
                        fx = (float)px;
                        fy = (float)py;
                        sc = BAND(py, px);
                    } else {
                        // score_box is a float32 view; box*grid promotes to float64 (grid is int64) and the 25
                        // products form a fresh contiguous array summed pairwise in float64; box.sum() reduces the
                        // strided float32 view row by row (inner runs of 2*rad+1 < 8 values, left to right, then
                        // accumulated across rows); box.mean() = that float32 sum / 25 in float32.
                        double prod_r[25 * 1], prod_c[25 * 1];
                        const int sidec = 2 * rad + 1;
                        float total = 0.0f;
                        int n = 0;
                        for (int q = -rad; q <= rad; q++) {
                            float rowsum = 0.0f;
                            for (int p = -rad; p <= rad; p++) {
                                const float v = BAND(py + q, px + p);
                                if (n < 25) {
                                    prod_r[n] = __dmul_rn((double)v, (double)q);  // x_grid = ROW offset (l.207 quirk)
                                    prod_c[n] = __dmul_rn((double)v, (double)p);  // y_grid = column offset
                                }
                                n++;
                                rowsum = __fadd_rn(rowsum, v);
                            }
                            total = (q == -rad) ? rowsum : __fadd_rn(total, rowsum);
                        }
                        const int cntp = min(n, 25);
                        const double denom = (double)total;
                        const double offx = __ddiv_rn(np_pairwise_sum_f64(prod_r, cntp), denom);
                        const double offy = __ddiv_rn(np_pairwise_sum_f64(prod_c, cntp), denom);
                        fx = __double2float_rn(__dadd_rn((double)px, offx));
                        fy = __double2float_rn(__dadd_rn((double)py, offy));
                        sc = __fdiv_rn(total, (float)(sidec * sidec));
                    }
                    DevPeak p;
                    p.x = (int)fx;  // truncated like the C++ route would (pafprocess.cpp:35-36)
                    p.y = (int)fy;
                    p.score = sc;
                    p.id = k;
                    out[k] = p;
                    if (a.subpix) {
                        float *sp = a.subpix + (((size_t)img * kParts + part) * a.cap + k) * 2;
                        sp[0] = fx;
                        sp[1] = fy;
                    }
                }
                __syncwarp();
            }
        }
#undef BAND
    }
    if (tid == 0) {
        a.part_count[img * kParts + part] = min(n_total, a.cap);
        if (n_total > a.cap) atomicOr(a.status + img, IBP_ST_PEAK_OVERFLOW);
    }
}

// After the per-part lists exist: running ids across parts (parse_skeletons.py:172-173) and the part-major "line"
// array (pafprocess.cpp:43-48) that results and grouping index by peak id.  One CTA per image.
__global__ void __launch_bounds__(128) pack_peaks_kernel(DevPeak *peaks, const int *part_count, int cap, ibp_peak *line, int line_stride,
                                                         int *n_peaks) {
    const int img = blockIdx.x;
    __shared__ int offs[kParts + 1];
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int p = 0; p < kParts; p++) {
            offs[p] = acc;
            acc += part_count[img * kParts + p];
        }
        offs[kParts] = acc;
        n_peaks[img] = min(acc, line_stride);
    }
    __syncthreads();
    for (int p = 0; p < kParts; p++) {
        const int n = part_count[img * kParts + p];
        DevPeak *src = peaks + ((size_t)img * kParts + p) * cap;
        for (int k = threadIdx.x; k < n; k += blockDim.x) {
            const int id = offs[p] + k;
            src[k].id = id;
            if (id < line_stride) {
                ibp_peak o;
                o.x = src[k].x;
                o.y = src[k].y;
                o.score = src[k].score;
                o.part = p;
                line[(size_t)img * line_stride + id] = o;
            }
        }
    }
}

}  // namespace ibp
