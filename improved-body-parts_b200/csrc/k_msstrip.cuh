// K1 (multi-scale), strip form — the same arithmetic as k_multiscale.cuh (utils/parse_skeletons.py:180-283, predict),
// organised so that nothing is computed twice and almost nothing goes through shared memory twice.
//
// One WARP owns a strip of 32 output columns of one channel of one image and walks DOWN the image, a segment of
// kSsRows output rows at a time, keeping for every scale of the search a rolling state instead of a tile:
//
//   feature rows ──A──► F (flip average, replicate border)           a few new rows per segment
//               ──B──► T1 rows (x4 resize, horizontal pass)         the last four rows stay in shared memory
//               ──C──► I_T[column][row group of 4]                  x4 resize, vertical pass; TRANSPOSED: the four phases of a
//                                                                   source row are contiguous, one 128-bit store
//               ──D──► H[4] = second resize, horizontal pass        a lane = an output column: 4 taps x one 128-bit load =
//                                                                   four rows at once, packed float32x2 over row pairs
//               ──E──► second resize, vertical pass, /n, float64    from the lane's sliding window of 8 H values in
//                                                                   REGISTERS (never stored); weights and tap positions of
//                                                                   an output row are warp-uniform and come from a table in
//                                                                   the kernel's parameter space (constant bank)
//
// Compared with the tile kernel: no vertical halo is recomputed (the walk is continuous), the second resize's
// intermediate image never exists, its vertical taps cost no shared-memory reads, and the per-row tap tables are built
// once on the host instead of once per tile and channel.  All synchronisation is __syncwarp.
//
// Replicate borders of the SECOND resize (it clamps to the cropped image [0,sh) x [0,sw)) are handled with "virtual"
// rows and columns: I(v, u) := I(clamp(v), clamp(u)), so the four taps of an output pixel are always consecutive.
// Virtual rows come in groups k = rows 4k+2 .. 4k+5 — exactly the four phases of feature row group k of the x4 resize.
#pragma once
#include "k_multiscale.cuh"

namespace ibp {

constexpr int kSsCols = 32;   // output columns of a strip = lanes
constexpr int kSsRows = 8;    // output rows of a segment (float64 accumulators in registers)
constexpr int kSsYCap = 960;  // entries of the per-row table in the parameter space (32 B each)
constexpr int kSsMaxNew = 8;  // feature rows fetched per phase
constexpr int kSsHist = 4;    // T1 rows kept between phases (three for the sliding window, one more for the bottom border)

struct __align__(16) SsXEntry {  // per output column and scale (global memory, read once per segment)
    int xv0;                     // first tap, virtual column (unclamped: sx - 1)
    int pad[3];
    float w[4];
};
struct __align__(16) SsYEntry {  // per output row and scale (parameter space)
    float w[4];
    int g;  // row group that completes this row's window: (y0 + 1) >> 2 with y0 = sy - 1
    int o;  // first tap's position in the window of 8: y0 - (4g - 2), 1..4
    int pad[2];
};
struct SsScale {
    const float *src;      // (batch, 2, 50, h, w)
    const SsXEntry *xtab;  // [W]
    int h, w, sh, sw;
    int s1;        // floats per T1 row
    int hist_off;  // float offset of this scale's history rows
    int kprime;    // last row group that holds a real row: (sh - 3) >> 2
    int cap;       // row groups one phase may produce
};
struct SsArgs {
    SsScale sc[kMsMaxScales];
    int n_scales, flip, H, W;
    float one;
    double *out;
    float *heat32;
    int y0, y1;       // output rows of this launch
    int chain_rows;   // rows one warp walks (multiple of kSsRows)
    int off_t1, off_hist, off_tw, off_pf;  // byte offsets into shared memory (I_T first; the F rows share its space)
    int f_stride;     // floats per F row
    SsYEntry ytab[kSsYCap];  // [scale][y1 - y0]
};

__device__ __forceinline__ void ms_planes_raw(const float *src, int h, int w, int img, int chan, const float *&upr, const float *&twin) {
    const size_t hw = (size_t)h * w;
    const float *base = src + (size_t)img * 2 * kLayers * hw;
    upr = base + (size_t)chan * hw;
    twin = base + (size_t)(kLayers + c_flip_src[chan]) * hw;
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ float2 ss_pair(const float *p) { return *reinterpret_cast<const float2 *>(p); }

// SLOTS = 16-byte slots (row groups) per column of I_T, odd so that the 128-bit stores of consecutive columns fall on
// different banks; a compile-time constant so that the tap addresses of stage D are immediates.
template <int SLOTS, bool POW2, bool TAIL>
__global__ void __launch_bounds__(kSsCols, 16) fuse_strip_kernel(const __grid_constant__ SsArgs a) {
    extern __shared__ __align__(16) unsigned char ss_smem[];
    constexpr int IS = SLOTS * 16;  // bytes per column of I_T
    const int lane = threadIdx.x;
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int X0 = blockIdx.x * kSsCols;
    const int Yc0 = a.y0 + blockIdx.y * a.chain_rows;
    const int Yc1 = min(Yc0 + a.chain_rows, a.y1);
    const int n_rows_tab = a.y1 - a.y0;
    const int Xl = min(X0 + lane, a.W - 1);
    const bool flip = a.flip != 0;

    unsigned char *IT = ss_smem;
    float *FS = reinterpret_cast<float *>(ss_smem);  // stage A's rows are consumed by B before C writes I_T
    float *T1S = reinterpret_cast<float *>(ss_smem + a.off_t1);
    float *HIST = reinterpret_cast<float *>(ss_smem + a.off_hist);
    float *TW = reinterpret_cast<float *>(ss_smem + a.off_tw);  // [4][4] phase weights, then the pairs (ph0,ph1) / (ph2,ph3) per tap
    int *ST = reinterpret_cast<int *>(TW + 32);                   // per scale: last group consumed, last T1 row produced
    int4 *GEO = reinterpret_cast<int4 *>(ST + 2 * kMsMaxScales);  // per scale: x_lo, n_c, kx_lo, n_kx of this strip
    float *FU = reinterpret_cast<float *>(ss_smem + a.off_pf);    // raw feature rows of the NEXT phase (upright, then twin), filled by cp.async
    float *FT = FU + kSsMaxNew * a.f_stride;
    bool pf_ready = false;  // the next produce() finds its first 32 feature columns in FU / FT

    if (lane < 4) {
        float w4[4];
        cubic_weights(__fadd_rn(__int2float_rn(lane) * 0.25f, 0.125f), w4);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            TW[lane * 4 + j] = w4[j];
            TW[16 + (lane >> 1) * 8 + j * 2 + (lane & 1)] = w4[j];
        }
    }
    if (lane < a.n_scales) {
        const SsScale &S = a.sc[lane];
        const int x_lo = __ldg(&S.xtab[X0].xv0);
        const int x_hi = __ldg(&S.xtab[min(X0 + kSsCols - 1, a.W - 1)].xv0) + 3;
        const int kx_lo = (clampi(x_lo, 0, S.sw - 1) - 2) >> 2, kx_hi = (clampi(x_hi, 0, S.sw - 1) - 2) >> 2;
        GEO[lane] = make_int4(x_lo, x_hi - x_lo + 1, kx_lo, kx_hi - kx_lo + 1);
    }
    __syncwarp();

    // the reference resizes the 30 limb channels and the 20 keypoint channels as two interleaved images: the last
    // (W*cn) % 4 elements of a destination row take the scalar tail of OpenCV's vertical pass (left-associated sum)
    const int cn = (chan < kLimbs) ? kLimbs : (kLayers - kLimbs);
    const int c_local = (chan < kLimbs) ? chan : chan - kLimbs;
    const bool my_tail = ((X0 + lane) * cn + c_local) >= ((a.W * cn) & ~3);
    const bool any_tail = __any_sync(0xffffffffu, my_tail);

    const float n_scales_f = __int2float_rn(a.n_scales);
    const float inv_scales = __fdiv_rn(1.0f, n_scales_f);
    const float2 one2 = dup2(a.one);

    float hst[kMsMaxScales][8];  // per scale: the lane's window (local memory; touched twice per phase)

    for (int Ys = Yc0; Ys < Yc1; Ys += kSsRows) {
        double acc[kSsRows];
#pragma unroll
        for (int j = 0; j < kSsRows; j++) acc[j] = 0.0;
        const int n_valid = min(kSsRows, Yc1 - Ys);

        for (int s = 0; s < a.n_scales; s++) {
            const SsScale &S = a.sc[s];
            const int yi0 = s * n_rows_tab + Ys - a.y0;  // a.ytab[yi0 + j] = row Ys + j (read by index: constant-bank loads)
            const int4 geo = GEO[s];
            const int x_lo = geo.x, n_c = geo.y, kx_lo = geo.z, n_kx = geo.w;
            const int4 xe_i = __ldg(reinterpret_cast<const int4 *>(&S.xtab[Xl]));
            const float4 xe_w = __ldg(reinterpret_cast<const float4 *>(&S.xtab[Xl].w[0]));
            const int s1 = S.s1;
            float *hist = HIST + S.hist_off;

            const int g_need = a.ytab[yi0 + n_valid - 1].g;
            int g_cur, t1_top;
            float2 P01, P23, H01, H23;  // window: rows 4g-2 .. 4g+5 of the lane's column after group g
            if (Ys == Yc0) {
                g_cur = a.ytab[yi0].g - 2;
                t1_top = g_cur - 1;
                P01 = P23 = H01 = H23 = make_float2(0.f, 0.f);
            } else {
                g_cur = ST[s];
                t1_top = ST[kMsMaxScales + s];
                P01 = make_float2(hst[s][0], hst[s][1]);
                P23 = make_float2(hst[s][2], hst[s][3]);
                H01 = make_float2(hst[s][4], hst[s][5]);
                H23 = make_float2(hst[s][6], hst[s][7]);
            }
            const float2 w0 = dup2(xe_w.x), w1 = dup2(xe_w.y), w2 = dup2(xe_w.z), w3 = dup2(xe_w.w);
            const unsigned char *d_base = IT + (xe_i.x - x_lo) * IS;

            // one group of four rows through stage D for the lane's column
            auto stage_d = [&](int slot) {
                const unsigned char *p = d_base + slot * 16;
                const float4 t0 = *reinterpret_cast<const float4 *>(p);
                const float4 t1 = *reinterpret_cast<const float4 *>(p + IS);
                const float4 t2 = *reinterpret_cast<const float4 *>(p + 2 * IS);
                const float4 t3 = *reinterpret_cast<const float4 *>(p + 3 * IS);
                P01 = H01;
                P23 = H23;
                H01 = hsum4x2(make_float2(t0.x, t0.y), make_float2(t1.x, t1.y), make_float2(t2.x, t2.y), make_float2(t3.x, t3.y), w0, w1, w2, w3, one2);
                H23 = hsum4x2(make_float2(t0.z, t0.w), make_float2(t1.z, t1.w), make_float2(t2.z, t2.w), make_float2(t3.z, t3.w), w0, w1, w2, w3, one2);
            };

            // stages A, B, C for row groups k1 .. k1+n_g-1 (into slots 0 ..)
            auto produce = [&](int k1, int n_g, bool staged) {
                const float *upr, *twin;
                ms_planes_raw(S.src, S.h, S.w, img, chan, upr, twin);
                const int n_new = k1 + n_g + 1 - t1_top;  // T1 rows t1_top+1 .. k1+n_g+1
                const int n_fx = n_kx + 3;
                const float2 a0 = ss_pair(TW + 16), a1 = ss_pair(TW + 18), a2 = ss_pair(TW + 20), a3 = ss_pair(TW + 22);
                const float2 b0 = ss_pair(TW + 24), b1 = ss_pair(TW + 26), b2 = ss_pair(TW + 28), b3 = ss_pair(TW + 30);
                // ---- A: feature rows, flip average on the way in (l.232-248), replicate border of the FIRST resize.  The
                //         first 32 columns were requested a phase ago (cp.async, see below) when `staged` ----
                if (staged) {
                    cp_async_wait_all();
                    __syncwarp();
                    if (lane < n_fx) {
#pragma unroll
                        for (int i = 0; i < kSsMaxNew; i++) {
                            if (i >= n_new) break;
                            const float u = FU[i * a.f_stride + lane];
                            FS[i * a.f_stride + lane] = flip ? avg_like_numpy(u, FT[i * a.f_stride + lane]) : u;
                        }
                    }
                }
                for (int c0 = staged ? kSsCols : 0; c0 < n_fx; c0 += kSsCols) {
                    const int c = c0 + lane;
                    const int fx = clampi(kx_lo - 1 + c, 0, S.w - 1);
                    const float *pu_col = upr + fx, *pt_col = twin + (S.w - 1 - fx);
                    float pu[kSsMaxNew], pt[kSsMaxNew];
#pragma unroll
                    for (int i = 0; i < kSsMaxNew; i++) {
                        if (i >= n_new) break;
                        const int fr = clampi(t1_top + 1 + i, 0, S.h - 1) * S.w;
                        pu[i] = __ldg(pu_col + fr);
                        pt[i] = flip ? __ldg(pt_col + fr) : 0.f;
                    }
                    if (c < n_fx) {
#pragma unroll
                        for (int i = 0; i < kSsMaxNew; i++) {
                            if (i >= n_new) break;
                            FS[i * a.f_stride + c] = flip ? avg_like_numpy(pu[i], pt[i]) : pu[i];
                        }
                    }
                }
                const bool border = (4 * k1 + 2 < 0) || (4 * (k1 + n_g - 1) + 5 > S.sh - 1);
                // the history rows in front of the new ones: T1S row r holds T1 row t1_top - 3 + r (only those C will read)
                {
                    const int r_lo = max(0, k1 + 2 - t1_top - (border ? 1 : 0));
                    for (int c = lane * 4; c < 4 * n_kx; c += 4 * kSsCols)
                        for (int r = r_lo; r < kSsHist; r++)
                            *reinterpret_cast<float4 *>(T1S + r * s1 + c) = *reinterpret_cast<const float4 *>(hist + ((t1_top - 3 + r + 4096) & 3) * s1 + c);
                }
                __syncwarp();
                // ---- B: first resize, horizontal pass: item = (new row, group of four columns) ----
                {
                    const float inv = __frcp_rn(__int2float_rn(n_kx));
                    for (int it = lane; it < n_new * n_kx; it += kSsCols) {
                        const int i = ms_div(it, inv), k = it - i * n_kx;
                        const float *row = FS + i * a.f_stride + k;
                        const float2 f0 = dup2(row[0]), f1 = dup2(row[1]), f2 = dup2(row[2]), f3 = dup2(row[3]);
                        const float2 lo = hsum4x2(f0, f1, f2, f3, a0, a1, a2, a3, one2);
                        const float2 hi = hsum4x2(f0, f1, f2, f3, b0, b1, b2, b3, one2);
                        *reinterpret_cast<float4 *>(T1S + (kSsHist + i) * s1 + 4 * k) = make_float4(lo.x, lo.y, hi.x, hi.y);
                    }
                }
                __syncwarp();
                // ---- C: first resize, vertical pass (always OpenCV's vector formula), into the transposed I_T ----
                {
                    const int t1_x0 = 4 * kx_lo + 2;
                    const float *row_km1 = T1S + (k1 - 1 - t1_top + 3) * s1;  // T1 row k1 - 1
                    for (int ci = lane; ci < n_c; ci += kSsCols) {
                        const int tc = clampi(x_lo + ci, 0, S.sw - 1) - t1_x0;
                        unsigned char *q = IT + ci * IS;
                        const float *tp = row_km1 + tc;
                        if (!border) {
                            float2 t0 = dup2(tp[0]), t1 = dup2(tp[s1]), t2 = dup2(tp[2 * s1]);
                            tp += 3 * s1;
                            for (int g = 0; g < n_g; g++) {
                                const float2 t3 = dup2(*tp);
                                const float2 lo = vsum4x2(t0, t1, t2, t3, a0, a1, a2, a3, one2);
                                const float2 hi = vsum4x2(t0, t1, t2, t3, b0, b1, b2, b3, one2);
                                *reinterpret_cast<float4 *>(q) = make_float4(lo.x, lo.y, hi.x, hi.y);
                                t0 = t1, t1 = t2, t2 = t3;
                                tp += s1;
                                q += 16;
                            }
                        } else {  // a group that touches the border of the cropped image: its rows clamp to real ones
                            for (int g = 0; g < n_g; g++) {
                                float r[4];
#pragma unroll
                                for (int p = 0; p < 4; p++) {
                                    const int ve = clampi(4 * (k1 + g) + 2 + p, 0, S.sh - 1);
                                    const int ke = (ve - 2) >> 2, pe = (ve - 2) & 3;
                                    const float *t = tp + (ke - k1) * s1;  // T1 row ke - 1 of this column
                                    r[p] = vsum4(t[0], t[s1], t[2 * s1], t[3 * s1], TW + 4 * pe);
                                }
                                *reinterpret_cast<float4 *>(q + g * 16) = make_float4(r[0], r[1], r[2], r[3]);
                            }
                        }
                    }
                }
                __syncwarp();
                // the last four T1 rows stay behind for the next phase (slot = row & 3; older ones already sit there)
                {
                    const int new_top = t1_top + n_new;
                    for (int kk = max(t1_top + 1, new_top - 3); kk <= new_top; kk++) {
                        const float *src = T1S + (kk - t1_top + 3) * s1;
                        float *dst = hist + ((kk + 4096) & 3) * s1;
                        for (int c = lane * 4; c < 4 * n_kx; c += 4 * kSsCols) *reinterpret_cast<float4 *>(dst + c) = *reinterpret_cast<const float4 *>(src + c);
                    }
                    t1_top = new_top;
                }
                __syncwarp();
            };

            int k1 = g_cur + 1;  // group in slot 0
            for (int pass = (Ys == Yc0) ? 0 : 1; pass < 2; pass++) {
                // pass 0 (start of the walk only): the two groups that fill the window up to the first row's
                const int n_g = pass == 0 ? 2 : g_need - g_cur;
                if (n_g > 0) {
                    produce(k1, n_g, pf_ready);
                    pf_ready = false;
                }
                if (pass == 0) {
                    stage_d(0);
                    stage_d(1);
                    g_cur += 2;
                    k1 = g_cur + 1;
                    __syncwarp();
                }
            }

            // ---- the feature rows of the NEXT phase set off now (cp.async), to land under this phase's D and E ----
            {
                int s2 = s + 1, Ys2 = Ys;
                if (s2 == a.n_scales) s2 = 0, Ys2 = Ys + kSsRows;
                if (Ys2 < Yc1) {
                    const SsScale &S2 = a.sc[s2];
                    const int yi2 = s2 * n_rows_tab + Ys2 - a.y0;
                    int gc2, tt2, ng2;
                    if (Ys2 == Yc0) {
                        gc2 = a.ytab[yi2].g - 2, tt2 = gc2 - 1, ng2 = 2;
                    } else {
                        gc2 = (s2 == s) ? g_need : ST[s2];
                        tt2 = (s2 == s) ? t1_top : ST[kMsMaxScales + s2];
                        ng2 = a.ytab[yi2 + min(kSsRows, Yc1 - Ys2) - 1].g - gc2;
                    }
                    if (ng2 > 0) {
                        const int n_new2 = gc2 + ng2 + 2 - tt2;
                        const int4 geo2 = GEO[s2];
                        if (lane < geo2.w + 3) {
                            const float *upr2, *twin2;
                            ms_planes_raw(S2.src, S2.h, S2.w, img, chan, upr2, twin2);
                            const int fx = clampi(geo2.z - 1 + lane, 0, S2.w - 1);
                            upr2 += fx;
                            twin2 += S2.w - 1 - fx;
#pragma unroll
                            for (int i = 0; i < kSsMaxNew; i++) {
                                if (i >= n_new2) break;
                                const int fr = clampi(tt2 + 1 + i, 0, S2.h - 1) * S2.w;
                                cp_async4(FU + i * a.f_stride + lane, upr2 + fr);
                                if (flip) cp_async4(FT + i * a.f_stride + lane, twin2 + fr);
                            }
                        }
                        cp_async_commit();
                        pf_ready = true;
                    }
                }
            }

            // ---- D + E for the rows of the segment ----
#pragma unroll
            for (int j = 0; j < kSsRows; j++) {
                if (j < n_valid) {
                    const float4 yw4 = *reinterpret_cast<const float4 *>(&a.ytab[yi0 + j].w[0]);
                    const int gj = a.ytab[yi0 + j].g, oj = a.ytab[yi0 + j].o;
                    const float yw[4] = {yw4.x, yw4.y, yw4.z, yw4.w};
                    while (g_cur < gj) {
                        g_cur++;
                        stage_d(g_cur - k1);
                    }
                    // the four taps sit at position oj of the window; one straight-line variant each, no register shuffling
                    float v;
                    switch (oj) {
                        case 1: v = vsum4(P01.y, P23.x, P23.y, H01.x, yw); break;
                        case 2: v = vsum4(P23.x, P23.y, H01.x, H01.y, yw); break;
                        case 3: v = vsum4(P23.y, H01.x, H01.y, H23.x, yw); break;
                        default: v = vsum4(H01.x, H01.y, H23.x, H23.y, yw); break;
                    }
                    if (TAIL && any_tail) {  // a strip at the right edge of an image whose rows do not end on a multiple of four elements
                        const float wv[8] = {P01.x, P01.y, P23.x, P23.y, H01.x, H01.y, H23.x, H23.y};
                        float t4[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) t4[q] = oj == 1 ? wv[1 + q] : (oj == 2 ? wv[2 + q] : (oj == 3 ? wv[3 + q] : wv[4 + q]));
                        if (my_tail) v = vsum4_tail(t4[0], t4[1], t4[2], t4[3], yw);
                    }
                    if (POW2) v = __fmul_rn(v, inv_scales);
                    else v = __fdiv_rn(v, n_scales_f);
                    acc[j] = __dadd_rn(acc[j], (double)v);
                }
            }
            ST[s] = g_cur;
            ST[kMsMaxScales + s] = t1_top;
            hst[s][0] = P01.x, hst[s][1] = P01.y, hst[s][2] = P23.x, hst[s][3] = P23.y;
            hst[s][4] = H01.x, hst[s][5] = H01.y, hst[s][6] = H23.x, hst[s][7] = H23.y;
            __syncwarp();  // I_T and the T1 scratch are free for the next scale
        }

        if (X0 + lane < a.W) {
            double *out = a.out + ((size_t)img * kLayers + chan) * (size_t)a.H * a.W;
            float *h32 = (a.heat32 && chan >= kLimbs && chan < kLimbs + kParts) ? a.heat32 + ((size_t)img * kParts + (chan - kLimbs)) * (size_t)a.H * a.W
                                                                                 : nullptr;
#pragma unroll
            for (int j = 0; j < kSsRows; j++) {
                if (j < n_valid) {
                    const size_t o = (size_t)(Ys + j) * a.W + (X0 + lane);
                    out[o] = acc[j];
                    if (h32) h32[o] = __double2float_rn(acc[j]);
                }
            }
        }
    }
}

}  // namespace ibp
