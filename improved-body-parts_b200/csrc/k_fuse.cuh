// K1 — flip / channel-permutation average of the network output.
// Reference: utils/parse_skeletons.py:80-96 (predict_refactor) == :232-248 (predict):
//     avg[:, :, c] = (out[0][c] + out[1][perm[c]][:, ::-1]) / 2
// computed in the model's dtype (fp16 under amp O1, l.75) and cast to float32 at the end (l.103).
//
// Streaming kernel, HBM-bound: per image it reads 2*50*h*w elements and writes 50*h*w floats (9.83 MB at 128x128
// fp32).  One thread produces kVec consecutive outputs from one 128-bit load of the upright plane and one 128-bit
// load of the mirrored plane (reversed in registers), one 128-bit store.
#pragma once
#include "ibp_device.cuh"

namespace ibp {

template <typename T>
struct FuseOps;

template <>
struct FuseOps<float> {
    static __device__ __forceinline__ float avg(float a, float b) { return __fmul_rn(__fadd_rn(a, b), 0.5f); }
    static __device__ __forceinline__ float one(float a) { return a; }
};
template <>
struct FuseOps<__half> {
    // numpy float16: each ufunc rounds its result to half — (a+b) -> half, then /2 -> half
    static __device__ __forceinline__ float avg(__half a, __half b) {
        const __half s = __float2half_rn(__fadd_rn(__half2float(a), __half2float(b)));
        return __half2float(__float2half_rn(__fmul_rn(__half2float(s), 0.5f)));
    }
    static __device__ __forceinline__ float one(__half a) { return __half2float(a); }
};

// vectorised path: w % 4 == 0 (fp32) so that rows are 16-byte aligned both forwards and mirrored
__global__ void __launch_bounds__(256) fuse_flip_f32_vec4(const float *__restrict__ in, float *__restrict__ out, int hw4, int w4, int flip) {
    // grid: x = vectors of one plane, y = channel, z = image
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t plane = (size_t)hw4 * 4;
    const float4 *up = reinterpret_cast<const float4 *>(in + ((size_t)b * 2 * kLayers + c) * plane);
    const float4 *fl = reinterpret_cast<const float4 *>(in + ((size_t)b * 2 * kLayers + kLayers + c_flip_src[c]) * plane);
    float4 *o = reinterpret_cast<float4 *>(out + ((size_t)b * kLayers + c) * plane);
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < hw4; v += gridDim.x * blockDim.x) {
        float4 a = __ldcs(up + v);
        if (flip) {
            const int row = v / w4, col = v - row * w4;
            const float4 m = __ldcs(fl + row * w4 + (w4 - 1 - col));
            a.x = FuseOps<float>::avg(a.x, m.w);
            a.y = FuseOps<float>::avg(a.y, m.z);
            a.z = FuseOps<float>::avg(a.z, m.y);
            a.w = FuseOps<float>::avg(a.w, m.x);
        }
        o[v] = a;
    }
}

// fp16 input, w % 8 == 0: 8 halves (one 128-bit load) per thread, two float4 stores
__global__ void __launch_bounds__(256) fuse_flip_f16_vec8(const __half *__restrict__ in, float *__restrict__ out, int hw8, int w8, int flip) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t plane = (size_t)hw8 * 8;
    const uint4 *up = reinterpret_cast<const uint4 *>(in + ((size_t)b * 2 * kLayers + c) * plane);
    const uint4 *fl = reinterpret_cast<const uint4 *>(in + ((size_t)b * 2 * kLayers + kLayers + c_flip_src[c]) * plane);
    float4 *o = reinterpret_cast<float4 *>(out + ((size_t)b * kLayers + c) * plane);
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < hw8; v += gridDim.x * blockDim.x) {
        union { uint4 u; __half h[8]; } a, m;
        a.u = __ldcs(up + v);
        float r[8];
        if (flip) {
            const int row = v / w8, col = v - row * w8;
            m.u = __ldcs(fl + row * w8 + (w8 - 1 - col));
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = FuseOps<__half>::avg(a.h[k], m.h[7 - k]);
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = __half2float(a.h[k]);
        }
        o[2 * v] = make_float4(r[0], r[1], r[2], r[3]);
        o[2 * v + 1] = make_float4(r[4], r[5], r[6], r[7]);
    }
}

// any width / alignment
template <typename T>
__global__ void __launch_bounds__(256) fuse_flip_scalar(const T *__restrict__ in, float *__restrict__ out, int h, int w, int flip) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t plane = (size_t)h * w;
    const T *up = in + ((size_t)b * 2 * kLayers + c) * plane;
    const T *fl = in + ((size_t)b * 2 * kLayers + kLayers + c_flip_src[c]) * plane;
    float *o = out + ((size_t)b * kLayers + c) * plane;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < (int)plane; i += gridDim.x * blockDim.x) {
        const int row = i / w, col = i - row * w;
        o[i] = flip ? FuseOps<T>::avg(up[i], fl[row * w + (w - 1 - col)]) : FuseOps<T>::one(up[i]);
    }
}

}  // namespace ibp
