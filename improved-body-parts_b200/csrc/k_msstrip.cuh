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
constexpr int kSsYCap = 1280; // entries of the per-row table in the parameter space (24 B each)
constexpr int kSsMaxNew = 8;  // feature rows fetched per phase (register prefetch depth of stage A)

struct __align__(16) SsXEntry {  // per output column and scale (global memory, read once per segment)
    int xv0;                     // first tap, virtual column (unclamped: sx - 1)
    int pad[3];
    float w[4];
};
struct SsYEntry {  // per output row and scale (parameter space)
    float w[4];
    int g;  // row group that completes this row's window: (y0 + 1) >> 2 with y0 = sy - 1
    int o;  // first tap's position in the window of 8: y0 - (4g - 2), 1..4
};
struct SsScale {
    const float *src;      // (batch, 2, 50, h, w)
    const SsXEntry *xtab;  // [W]
    int h, w, sh, sw;
    int s1;        // floats per T1 row
    int hist_off;  // float offset of this scale's four history rows
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
    int it_stride;    // bytes per column of I_T (row groups x 16)
    int off_t1, off_f, off_hist, off_tw;  // byte offsets into shared memory (I_T first)
    int f_stride;     // floats per F row
    SsYEntry ytab[kSsYCap];  // [scale][y1 - y0]
};

__device__ __forceinline__ void ms_planes_raw(const float *src, int h, int w, int img, int chan, const float *&upr, const float *&twin) {
    const size_t hw = (size_t)h * w;
    const float *base = src + (size_t)img * 2 * kLayers * hw;
    upr = base + (size_t)chan * hw;
    twin = base + (size_t)(kLayers + c_flip_src[chan]) * hw;
}
__device__ __forceinline__ float2 ss_pair(const float *p) { return *reinterpret_cast<const float2 *>(p); }

template <bool POW2>
__global__ void __launch_bounds__(kSsCols, 20) fuse_strip_kernel(const __grid_constant__ SsArgs a) {
    extern __shared__ __align__(16) unsigned char ss_smem[];
    const int lane = threadIdx.x;
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int X0 = blockIdx.x * kSsCols;
    const int Yc0 = a.y0 + blockIdx.y * a.chain_rows;
    const int Yc1 = min(Yc0 + a.chain_rows, a.y1);
    const int n_rows_tab = a.y1 - a.y0;
    const int Xl = min(X0 + lane, a.W - 1);
    const bool flip = a.flip != 0;

    unsigned char *IT = ss_smem;
    float *T1S = reinterpret_cast<float *>(ss_smem + a.off_t1);
    float *FS = reinterpret_cast<float *>(ss_smem + a.off_f);
    float *HIST = reinterpret_cast<float *>(ss_smem + a.off_hist);
    float *TW = reinterpret_cast<float *>(ss_smem + a.off_tw);  // [4][4] phase weights, then the pairs (ph0,ph1) / (ph2,ph3) per tap
    int *ST = reinterpret_cast<int *>(TW + 32);                   // per scale: last group consumed, last T1 row produced

    if (lane < 4) {
        float w4[4];
        cubic_weights(__fadd_rn(__int2float_rn(lane) * 0.25f, 0.125f), w4);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            TW[lane * 4 + j] = w4[j];
            TW[16 + (lane >> 1) * 8 + j * 2 + (lane & 1)] = w4[j];
        }
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
            // ---- strip geometry of this scale (warp-uniform) and the lane's own column ----
            const int x_lo = __ldg(&S.xtab[X0].xv0);
            const int x_hi = __ldg(&S.xtab[min(X0 + kSsCols - 1, a.W - 1)].xv0) + 3;
            const int4 xe_i = __ldg(reinterpret_cast<const int4 *>(&S.xtab[Xl]));
            const float4 xe_w = __ldg(reinterpret_cast<const float4 *>(&S.xtab[Xl].w[0]));
            const int n_c = x_hi - x_lo + 1;
            const int kx_lo = (clampi(x_lo, 0, S.sw - 1) - 2) >> 2, kx_hi = (clampi(x_hi, 0, S.sw - 1) - 2) >> 2;
            const int n_kx = kx_hi - kx_lo + 1, n_fx = n_kx + 3, t1_x0 = 4 * kx_lo + 2;
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
            const unsigned char *d_base = IT + (xe_i.x - x_lo) * a.it_stride;

            // one group of four rows through stage D for the lane's column
            auto stage_d = [&](int slot) {
                const unsigned char *p = d_base + slot * 16;
                const float4 t0 = *reinterpret_cast<const float4 *>(p);
                const float4 t1 = *reinterpret_cast<const float4 *>(p + a.it_stride);
                const float4 t2 = *reinterpret_cast<const float4 *>(p + 2 * a.it_stride);
                const float4 t3 = *reinterpret_cast<const float4 *>(p + 3 * a.it_stride);
                P01 = H01;
                P23 = H23;
                H01 = hsum4x2(make_float2(t0.x, t0.y), make_float2(t1.x, t1.y), make_float2(t2.x, t2.y), make_float2(t3.x, t3.y), w0, w1, w2, w3, one2);
                H23 = hsum4x2(make_float2(t0.z, t0.w), make_float2(t1.z, t1.w), make_float2(t2.z, t2.w), make_float2(t3.z, t3.w), w0, w1, w2, w3, one2);
            };

            // stages A, B, C for row groups k1 .. k1+n_g-1 (into slots 0 ..)
            auto produce = [&](int k1, int n_g) {
                const float *upr, *twin;
                ms_planes_raw(S.src, S.h, S.w, img, chan, upr, twin);
                const int n_new = k1 + n_g + 1 - t1_top;  // T1 rows t1_top+1 .. k1+n_g+1
                // ---- A: feature rows, flip average on the way in (l.232-248), replicate border of the FIRST resize ----
                for (int c0 = 0; c0 < n_fx; c0 += kSsCols) {
                    const int c = c0 + lane;
                    const int fx = clampi(kx_lo - 1 + c, 0, S.w - 1);
                    float pu[kSsMaxNew], pt[kSsMaxNew];
#pragma unroll
                    for (int i = 0; i < kSsMaxNew; i++) {
                        if (i < n_new && c < n_fx) {
                            const int fr = clampi(t1_top + 1 + i, 0, S.h - 1);
                            pu[i] = __ldg(upr + fr * S.w + fx);
                            if (flip) pt[i] = __ldg(twin + fr * S.w + (S.w - 1 - fx));
                        }
                    }
#pragma unroll
                    for (int i = 0; i < kSsMaxNew; i++)
                        if (i < n_new && c < n_fx) FS[i * a.f_stride + c] = flip ? avg_like_numpy(pu[i], pt[i]) : pu[i];
                }
                __syncwarp();
                // ---- B: first resize, horizontal pass: item = (new row, group of four columns) ----
                {
                    const float2 a0 = ss_pair(TW + 16), a1 = ss_pair(TW + 18), a2 = ss_pair(TW + 20), a3 = ss_pair(TW + 22);
                    const float2 b0 = ss_pair(TW + 24), b1 = ss_pair(TW + 26), b2 = ss_pair(TW + 28), b3 = ss_pair(TW + 30);
                    const float inv = __frcp_rn(__int2float_rn(n_kx));
                    for (int it = lane; it < n_new * n_kx; it += kSsCols) {
                        const int i = ms_div(it, inv), k = it - i * n_kx;
                        const float *row = FS + i * a.f_stride + k;
                        const float2 f0 = dup2(row[0]), f1 = dup2(row[1]), f2 = dup2(row[2]), f3 = dup2(row[3]);
                        const float2 lo = hsum4x2(f0, f1, f2, f3, a0, a1, a2, a3, one2);
                        const float2 hi = hsum4x2(f0, f1, f2, f3, b0, b1, b2, b3, one2);
                        *reinterpret_cast<float4 *>(T1S + i * s1 + 4 * k) = make_float4(lo.x, lo.y, hi.x, hi.y);
                    }
                }
                __syncwarp();
                // ---- C: first resize, vertical pass (always OpenCV's vector formula), into the transposed I_T ----
                {
                    const float2 a0 = ss_pair(TW + 16), a1 = ss_pair(TW + 18), a2 = ss_pair(TW + 20), a3 = ss_pair(TW + 22);
                    const float2 b0 = ss_pair(TW + 24), b1 = ss_pair(TW + 26), b2 = ss_pair(TW + 28), b3 = ss_pair(TW + 30);
                    // T1 row kk lives in the scratch when it is new, else in its history slot
                    auto t1row = [&](int kk) -> const float * { return kk > t1_top ? T1S + (kk - t1_top - 1) * s1 : hist + ((kk + 4096) & 3) * s1; };
                    for (int ci = lane; ci < n_c; ci += kSsCols) {
                        const int tc = clampi(x_lo + ci, 0, S.sw - 1) - t1_x0;
                        unsigned char *q = IT + ci * a.it_stride;
                        float2 t0 = dup2(t1row(k1 - 1)[tc]), t1 = dup2(t1row(k1)[tc]), t2 = dup2(t1row(k1 + 1)[tc]);
                        for (int g = 0; g < n_g; g++) {
                            const int k = k1 + g;
                            const float2 t3 = dup2(t1row(k + 2)[tc]);
                            float4 v;
                            if (4 * k + 2 >= 0 && 4 * k + 5 <= S.sh - 1) {
                                const float2 lo = vsum4x2(t0, t1, t2, t3, a0, a1, a2, a3, one2);
                                const float2 hi = vsum4x2(t0, t1, t2, t3, b0, b1, b2, b3, one2);
                                v = make_float4(lo.x, lo.y, hi.x, hi.y);
                            } else {  // a group that touches the border of the cropped image: rows clamp to real ones
                                float r[4];
#pragma unroll
                                for (int p = 0; p < 4; p++) {
                                    const int ve = clampi(4 * k + 2 + p, 0, S.sh - 1);
                                    const int ke = (ve - 2) >> 2, pe = (ve - 2) & 3;
                                    r[p] = vsum4(t1row(ke - 1)[tc], t1row(ke)[tc], t1row(ke + 1)[tc], t1row(ke + 2)[tc], TW + 4 * pe);
                                }
                                v = make_float4(r[0], r[1], r[2], r[3]);
                            }
                            *reinterpret_cast<float4 *>(q + g * 16) = v;
                            t0 = t1, t1 = t2, t2 = t3;
                        }
                    }
                }
                __syncwarp();
                // the last four T1 rows stay behind for the next phase (slot = row & 3; older ones already sit there)
                {
                    const int new_top = t1_top + n_new;
                    for (int kk = max(t1_top + 1, new_top - 3); kk <= new_top; kk++) {
                        const float *src = T1S + (kk - t1_top - 1) * s1;
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
                if (n_g > 0) produce(k1, n_g);
                if (pass == 0) {
                    stage_d(0);
                    stage_d(1);
                    g_cur += 2;
                    k1 = g_cur + 1;
                    __syncwarp();
                }
            }

            // ---- D + E for the rows of the segment ----
#pragma unroll
            for (int j = 0; j < kSsRows; j++) {
                if (j < n_valid) {
                    const int gj = a.ytab[yi0 + j].g, oj = a.ytab[yi0 + j].o;
                    const float yw[4] = {a.ytab[yi0 + j].w[0], a.ytab[yi0 + j].w[1], a.ytab[yi0 + j].w[2], a.ytab[yi0 + j].w[3]};
                    while (g_cur < gj) {
                        g_cur++;
                        stage_d(g_cur - k1);
                    }
                    float s0, s1v, s2, s3;
                    switch (oj) {
                        case 1: s0 = P01.y, s1v = P23.x, s2 = P23.y, s3 = H01.x; break;
                        case 2: s0 = P23.x, s1v = P23.y, s2 = H01.x, s3 = H01.y; break;
                        case 3: s0 = P23.y, s1v = H01.x, s2 = H01.y, s3 = H23.x; break;
                        default: s0 = H01.x, s1v = H01.y, s2 = H23.x, s3 = H23.y; break;
                    }
                    float v = vsum4(s0, s1v, s2, s3, yw);
                    if (any_tail && my_tail) v = vsum4_tail(s0, s1v, s2, s3, yw);
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
