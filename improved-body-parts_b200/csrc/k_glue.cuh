// The two ends of the hot path.
//
//   prepare_image_kernel   utils/parse_skeletons.py:52-72 after the cv2.resize: pad right/down to a multiple of
//                          `max_downsample` with `padValue` (utils/util.py:44-65), np.float32(img / 255) (a float64
//                          division, then one rounding to float32), the horizontally flipped twin next to it.
//   coco_rows_kernel       evaluate.py:182-209 append_result: CMU-18 -> COCO-17 by ORDER_COCO (evaluate.py:40),
//                          (x, y, 1) for a found part, zeros otherwise, the human's score; rows of all images compacted
//                          so that one small copy brings the whole batch to the host.
#pragma once
#include "ibp_device.cuh"

namespace ibp {

constexpr int kCocoParts = 17;
constexpr int kCocoRow = kCocoParts * 3 + 1;  // 17 x (x, y, v) + score
__constant__ int c_order_coco[kCocoParts] = {0, 15, 14, 17, 16, 5, 2, 6, 3, 7, 4, 11, 8, 12, 9, 13, 10};  // evaluate.py:40

struct PrepArgs {
    const uint8_t *img;  // (batch, h, w, 3) uint8 HWC (OpenCV BGR)
    float *out;          // (batch, flip ? 2 : 1, hp, wp, 3) float32 HWC
    int batch, h, w, hp, wp, flip;
    int pad_value;
};

// one thread per padded pixel: three values, written to the upright image and mirrored into the twin
__global__ void __launch_bounds__(256) prepare_image_kernel(PrepArgs a) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= a.wp) return;
    int v[3] = {a.pad_value, a.pad_value, a.pad_value};
    if (y < a.h && x < a.w) {
        const uint8_t *p = a.img + ((size_t)(b * a.h + y) * a.w + x) * 3;
        v[0] = p[0], v[1] = p[1], v[2] = p[2];
    }
    float f[3];
#pragma unroll
    for (int c = 0; c < 3; c++) f[c] = __double2float_rn(__ddiv_rn((double)v[c], 255.0));
    const size_t plane = (size_t)a.hp * a.wp * 3;
    float *o = a.out + (size_t)b * (a.flip ? 2 : 1) * plane + ((size_t)y * a.wp + x) * 3;
    o[0] = f[0], o[1] = f[1], o[2] = f[2];
    if (a.flip) {
        float *m = a.out + ((size_t)b * 2 + 1) * plane + ((size_t)y * a.wp + (a.wp - 1 - x)) * 3;
        m[0] = f[0], m[1] = f[1], m[2] = f[2];
    }
}

struct CocoArgs {
    const int *n_humans;      // [batch]
    const ibp_human *humans;  // [batch][human_stride]
    const ibp_peak *line;     // [batch][line_stride], index == peak id
    int batch, human_stride, line_stride;
    int *offsets;             // [batch + 1] first row of every image, total at the end
    float *rows;              // [total][kCocoRow]
};

// one CTA per image; its first row = the number of humans of the images before it
__global__ void __launch_bounds__(128) coco_rows_kernel(CocoArgs a) {
    const int b = blockIdx.x, tid = threadIdx.x;
    __shared__ int s_part[4];
    int before = 0;
    for (int i = tid; i < b; i += blockDim.x) before += min(a.n_humans[i], a.human_stride);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) before += __shfl_xor_sync(0xffffffffu, before, d);
    if ((tid & 31) == 0) s_part[tid >> 5] = before;
    __syncthreads();
    const int first = s_part[0] + s_part[1] + s_part[2] + s_part[3];
    const int n = min(a.n_humans[b], a.human_stride);
    if (tid == 0) {
        a.offsets[b] = first;
        if (b == a.batch - 1) a.offsets[a.batch] = first + n;
    }
    const ibp_human *hum = a.humans + (size_t)b * a.human_stride;
    const ibp_peak *line = a.line + (size_t)b * a.line_stride;
    for (int i = tid; i < n * kCocoParts; i += blockDim.x) {
        const int k = i / kCocoParts, j = i - k * kCocoParts;
        const int pid = hum[k].peak_id[c_order_coco[j]];
        float *row = a.rows + (size_t)(first + k) * kCocoRow;
        float x = 0.f, y = 0.f, v = 0.f;
        if (pid >= 0) {
            x = __int2float_rn(line[pid].x);
            y = __int2float_rn(line[pid].y);
            v = 1.f;
        }
        row[3 * j] = x, row[3 * j + 1] = y, row[3 * j + 2] = v;
        if (j == 0) row[kCocoRow - 1] = hum[k].score;
    }
}

}  // namespace ibp
