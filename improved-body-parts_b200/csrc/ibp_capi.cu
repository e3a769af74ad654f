// C ABI of libibp_b200.so (declared in include/ibp.h): handle management, launch wrappers, the legacy
// seven-call protocol of the reference's pafprocess module, and the batched pipeline.
// Host-side only: buffers, launch geometry, error reporting.  No compute happens on the CPU.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ibp_device.cuh"
#include "k_fuse.cuh"
#include "k_glue.cuh"
#include "k_limbs.cuh"
#include "k_limbs3.cuh"
#include "k_multiscale.cuh"
#include "k_msstrip.cuh"
#include "k_linear.cuh"
#include "k_peaks.cuh"
#include "k_pyroute.cuh"
#include "k_resize.cuh"
#include "k_warp.cuh"
#include "k_upsample.cuh"

using namespace ibp;

// ---------------------------------------------------------------------------------------------------------------
// Scratch of ibp_find_peaks_noalloc (include/ibp.h §5): ONE file-scope device array, held by every GPU that loads the
// library until the process exits.  In host code `g_noalloc_scratch` is a host-side stand-in, NOT a device address: the
// entry point asks the runtime for the device address once per device (cudaGetSymbolAddress, with that device current)
// and hands that pointer to the kernel.
// ---------------------------------------------------------------------------------------------------------------
struct NoAllocScratch {
    DevPeak band_peaks[IBP_NOALLOC_MAX_BATCH * kParts * IBP_NOALLOC_MAX_BANDS * IBP_NOALLOC_CAP];  // per-band lists
    DevPeak peaks[IBP_NOALLOC_MAX_BATCH * kParts * IBP_NOALLOC_CAP];                               // per-part lists
    // ---- everything from here on is cleared at the start of every call ----
    int band_count[IBP_NOALLOC_MAX_BATCH * kParts * IBP_NOALLOC_MAX_BANDS];
    unsigned ticket[IBP_NOALLOC_MAX_BATCH];
    unsigned batch_ticket;
    int order[IBP_NOALLOC_MAX_BATCH];
};
static_assert(sizeof(NoAllocScratch) <= IBP_NOALLOC_SCRATCH_BYTES, "IBP_NOALLOC_SCRATCH_BYTES too small");
static __device__ __align__(256) unsigned char g_noalloc_scratch[IBP_NOALLOC_SCRATCH_BYTES];

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e__ = (call);                                                                         \
        if (e__ != cudaSuccess) return fail(IBP_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

constexpr size_t kMaxDynSmem = 220 * 1024;  // opt-in dynamic shared memory we ask for (device max 227 KB minus static use)
constexpr size_t kPeakBandTarget = 16 * 1024;  // owned rows of one band (the CTA also stages 2*halo rows around them)
constexpr size_t kPeakBandTargetStream = 24 * 1024;
constexpr size_t kLimbPlaneMaxBytes = 96 * 1024;
constexpr int kL3MaxCtasPerSm = 4;  // limbs_stream3_kernel: CTAs per SM its global scratch is sized for

// bytes of one staged band: the streaming kernel (two buffer sets per CTA, few fat items) likes them larger than the
// one-CTA-per-band kernel (five CTAs per SM)
size_t peak_band_target(bool streaming) {
    static size_t v = 0;
    if (!v) {
        const char *e = getenv("IBP_PEAK_BAND_BYTES");  // tuning knob for experiments
        v = e ? (size_t)atol(e) : 1;
        if (e && v < 1024) v = 1024;
    }
    return v > 1 ? v : (streaming ? kPeakBandTargetStream : kPeakBandTarget);
}

template <typename T>
void free_dev(T *&p) {
    if (p) cudaFree(p);
    p = nullptr;
}
template <typename T>
void free_host(T *&p) {
    if (p) cudaFreeHost(p);
    p = nullptr;
}

}  // namespace

struct ibp_handle {
    int device = 0;
    int sm_count = 0;
    // streaming peak kernel: occupancy cached per (instantiation, shared bytes)
    const void *pk_occ_kern = nullptr;
    size_t pk_occ_smem = 0;
    int pk_occ = 0;
    unsigned *d_pk_work = nullptr;  // work counter of the streaming peak kernel, never reset ...
    unsigned pk_work_base = 0;      // ... its value before the next launch
    ibp_params P;
    ibp_limits L;
    int line_stride = 0;
    // device
    DevPeak *d_peaks = nullptr;
    double *d_subpix = nullptr;
    int *d_part_count = nullptr;
    DevLink *d_links = nullptr;
    int *d_link_count = nullptr;
    ibp_peak *d_line = nullptr;
    int *d_n_peaks = nullptr;
    ibp_human *d_humans = nullptr;
    int *d_n_humans = nullptr;
    uint32_t *d_status = nullptr;
    // scratch of the banded peak kernel and the two ticket counters (last CTA of an image packs / groups)
    DevPeak *d_band_peaks = nullptr;
    double *d_band_subpix = nullptr;
    int *d_band_count = nullptr;
    int band_slots = 0;  // nbands the scratch was sized for
    unsigned *d_ticket = nullptr;  // [2][max_batch] + 1
    // Python-semantics route (find_connections / find_humans), allocated on first use
    double *d_score64 = nullptr;      // (B,18,cap) float64 peak scores of ibp_set_peaks64
    int use_score64 = 0;
    double *d_py_conn = nullptr;      // (B,30,cap,6)
    int *d_py_conn_count = nullptr;   // (B,30)
    double *d_py_cand = nullptr;      // (B,line_stride)
    double *d_py_people = nullptr;    // (B,max_humans,20,2)
    int *d_py_n_people = nullptr;     // (B)
    int *d_order = nullptr;        // [max_batch] images by descending peak count (written by the peak kernel)
    int order_batch = 0;           // batch size d_order is valid for (0: not valid)
    int defer_bands = 0;           // > 0: the last peak launch left its band lists unpacked (this many bands per plane)
    int defer_batch = 0;
    int *d_yext = nullptr;         // [max_batch][18][4] y and x extent of each part's peaks, kept "empty" between launches
    int4 *d_item_meta = nullptr;   // [max_batch*30] limb items in heaviest-image-first order (index_bands_kernel)
    unsigned *d_lb_work = nullptr; // work counter of the streaming limb kernel, never reset ...
    unsigned lb_work_base = 0;     // ... its value before the next launch
    const void *lb_occ_kern = nullptr;
    size_t lb_occ_smem = 0;        // occupancy of the streaming limb kernel, cached per (instantiation, shared-memory size)
    int lb_occ = 0;
    const void *l3_occ_kern = nullptr;  // the same for limbs_stream3_kernel
    size_t l3_occ_smem = 0;
    int l3_occ = 0;
    unsigned char *d_l3_scratch = nullptr;  // per-CTA global scratch of limbs_stream3_kernel (kL3MaxCtasPerSm CTAs per SM)
    int *d_partbase = nullptr;     // [max_batch][19] first id of each part + image total (index_bands_kernel)
    int *d_n_raw = nullptr;        // [max_batch] peaks per image, accumulated by the streaming peak kernel (lives after d_status)
    unsigned long long *d_timeline = nullptr;  // experiments only (IBP_DEBUG & 8)
    unsigned long long *d_timeline_pk = nullptr;  // experiments only (IBP_DEBUG & 16)
    size_t timeline_pk_n = 0;
    float *d_fused = nullptr;
    size_t fused_floats = 0;
    // strip form of the multi-scale kernel: per-column tap tables of every scale (device + pinned staging), cached by shape
    SsXEntry *d_ms_xtab = nullptr, *h_ms_xtab = nullptr;
    size_t ms_xtab_cap = 0;
    std::vector<int> ms_xtab_key;
    // bilinear operator mode: composed tap tables of every scale (x then y), cached by shape
    LinTap *d_lin_tab = nullptr, *h_lin_tab = nullptr;
    size_t lin_tab_cap = 0;
    std::vector<int> lin_tab_key;
    void *d_stage = nullptr;  // host input staging / legacy limb map
    size_t stage_bytes = 0;
    // pinned host mirrors
    int *h_n_peaks = nullptr;
    int *h_part_count = nullptr;
    ibp_peak *h_line = nullptr;
    int *h_n_humans = nullptr;
    ibp_human *h_humans = nullptr;
    uint32_t *h_status = nullptr;
    DevPeak *h_peaks_stage = nullptr;  // one image worth of per-part lists for ibp_set_peaks
    // compacted COCO rows (ibp_fetch_coco), allocated on first use
    int *d_coco_off = nullptr, *h_coco_off = nullptr;
    float *d_coco_rows = nullptr, *h_coco_rows = nullptr;
    int64_t launches = 0;
    // optional per-kernel timing with events recorded on the launching stream (ibp_set_profiling)
    int profiling = 0;
    cudaEvent_t ev[2 * IBP_NUM_STAGES] = {nullptr};
    int ev_used[IBP_NUM_STAGES] = {0};
};

namespace {

// `st` is the stream the caller is about to copy into the buffer on: the zero fill is ordered on that same stream (a
// plain cudaMemset runs on the legacy default stream, which a cudaStreamNonBlocking stream never waits for)
int ensure_stage(ibp_handle *h, size_t bytes, cudaStream_t st) {
    if (h->stage_bytes >= bytes) return IBP_OK;
    if (h->d_stage) CU(cudaDeviceSynchronize());  // a regrow: nothing in flight may still read the old buffer
    free_dev(h->d_stage);
    h->stage_bytes = 0;
    CU(cudaMalloc(&h->d_stage, bytes));
    CU(cudaMemsetAsync(h->d_stage, 0, bytes, st));  // planes that are never uploaded (the background channels) stay zero
    h->stage_bytes = bytes;
    return IBP_OK;
}
int ensure_bands(ibp_handle *h, int nbands) {
    if (h->band_slots >= nbands) return IBP_OK;
    free_dev(h->d_band_peaks);
    free_dev(h->d_band_subpix);
    free_dev(h->d_band_count);
    h->band_slots = 0;
    const size_t n = (size_t)h->L.max_batch * kParts * nbands;
    CU(cudaMalloc(&h->d_band_peaks, n * h->L.max_peaks_per_part * sizeof(DevPeak)));
    CU(cudaMalloc(&h->d_band_subpix, n * h->L.max_peaks_per_part * 2 * sizeof(double)));
    CU(cudaMalloc(&h->d_band_count, n * sizeof(int)));
    h->band_slots = nbands;
    return IBP_OK;
}
int ensure_py(ibp_handle *h) {
    if (h->d_py_conn) return IBP_OK;
    const size_t B = h->L.max_batch, cap = h->L.max_peaks_per_part;
    CU(cudaMalloc(&h->d_score64, B * kParts * cap * sizeof(double)));
    CU(cudaMalloc(&h->d_py_conn, B * kLimbs * cap * 6 * sizeof(double)));
    CU(cudaMalloc(&h->d_py_conn_count, B * kLimbs * sizeof(int)));
    CU(cudaMalloc(&h->d_py_cand, B * h->line_stride * sizeof(double)));
    CU(cudaMalloc(&h->d_py_people, B * h->L.max_humans * 40 * sizeof(double)));
    CU(cudaMalloc(&h->d_py_n_people, B * sizeof(int)));
    CU(cudaMemset(h->d_py_conn_count, 0, B * kLimbs * sizeof(int)));
    CU(cudaMemset(h->d_py_n_people, 0, B * sizeof(int)));
    CU(cudaFuncSetAttribute(py_connections_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
    CU(cudaFuncSetAttribute(py_humans_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
    CU(cudaDeviceSynchronize());  // the fills above ran on the legacy stream: done before any user stream touches the buffers
    return IBP_OK;
}
int ensure_fused(ibp_handle *h, size_t floats) {
    if (h->fused_floats >= floats) return IBP_OK;
    free_dev(h->d_fused);
    h->fused_floats = 0;
    CU(cudaMalloc(&h->d_fused, floats * sizeof(float)));
    h->fused_floats = floats;
    return IBP_OK;
}

void destroy_buffers(ibp_handle *h) {
    free_dev(h->d_peaks);
    free_dev(h->d_subpix);
    free_dev(h->d_part_count);
    free_dev(h->d_links);
    free_dev(h->d_link_count);
    free_dev(h->d_line);
    free_dev(h->d_n_peaks);
    free_dev(h->d_humans);
    free_dev(h->d_n_humans);
    free_dev(h->d_status);
    free_dev(h->d_fused);
    free_dev(h->d_ms_xtab);
    free_host(h->h_ms_xtab);
    free_dev(h->d_lin_tab);
    free_host(h->h_lin_tab);
    free_dev(h->d_stage);
    free_dev(h->d_band_peaks);
    free_dev(h->d_band_subpix);
    free_dev(h->d_band_count);
    free_dev(h->d_ticket);
    free_dev(h->d_timeline);
    free_dev(h->d_timeline_pk);
    free_dev(h->d_order);
    free_dev(h->d_pk_work);
    free_dev(h->d_partbase);
    free_dev(h->d_item_meta);
    free_dev(h->d_yext);
    free_dev(h->d_lb_work);
    free_dev(h->d_l3_scratch);
    free_dev(h->d_score64);
    free_dev(h->d_py_conn);
    free_dev(h->d_py_conn_count);
    free_dev(h->d_py_cand);
    free_dev(h->d_py_people);
    free_dev(h->d_py_n_people);
    free_host(h->h_n_peaks);
    free_host(h->h_part_count);
    free_dev(h->d_coco_off);
    free_dev(h->d_coco_rows);
    free_host(h->h_coco_off);
    free_host(h->h_coco_rows);
    free_host(h->h_line);
    free_host(h->h_n_humans);
    free_host(h->h_humans);
    free_host(h->h_status);
    free_host(h->h_peaks_stage);
    for (int i = 0; i < 2 * IBP_NUM_STAGES; i++)
        if (h->ev[i]) {
            cudaEventDestroy(h->ev[i]);
            h->ev[i] = nullptr;
        }
}

// shared memory of the limb phase (layout in k_limbs.cuh) ...
size_t limb_phase_bytes(const ibp_handle *h, size_t plane_bytes) {
    size_t region0 = std::max(plane_bytes, (size_t)h->L.max_candidates_per_limb * 8);
    region0 = (region0 + 15) & ~(size_t)15;
    return region0 + (size_t)h->L.max_candidates_per_limb * 16 + (size_t)h->L.max_peaks_per_part * (2 * sizeof(DevPeak) + 2 + 8) + 32 +
           (size_t)kPairChunk * (sizeof(PairGeo) + (size_t)h->P.max_steps * 4) + 16;
}
// ... and of the grouping phase that the last CTA of an image runs in the same allocation (k_group.cuh): the
// skeleton table plus room for at least 256 staged connections
size_t group_phase_bytes(const ibp_handle *h) { return (size_t)h->L.max_bodies * kBodyBytes + 128 + 256 * sizeof(DevLink); }
size_t limb_smem_bytes(const ibp_handle *h, size_t plane_bytes) {
    return (std::max(limb_phase_bytes(h, plane_bytes), group_phase_bytes(h)) + 15) & ~(size_t)15;
}

int validate_limits(const ibp_limits &l) {
    if (l.max_batch < 1 || l.max_peaks_per_part < 1 || l.max_peaks_per_part > 32767 || l.max_candidates_per_limb < 32 ||
        (l.max_candidates_per_limb & (l.max_candidates_per_limb - 1)) != 0 || l.max_bodies < 32 || l.max_humans < 1 ||
        l.max_feat_h < 1 || l.max_feat_w < 1)
        return fail(IBP_ERR_ARG, "ibp_limits out of range (max_candidates_per_limb must be a power of two >= 32)");
    return IBP_OK;
}

// Launch `kern` as a programmatic dependent of the previous kernel in the stream: its CTAs may be set up while that
// kernel drains; the kernel itself waits (griddepcontrol.wait) before it touches anything the predecessor wrote.
template <typename Args>
cudaError_t launch_dependent(void (*kern)(Args), dim3 grid, dim3 block, size_t smem, cudaStream_t st, const Args &a) {
    static const bool off = getenv("IBP_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = off ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kern, a);
}

struct StageTimer {
    ibp_handle *h;
    int stage;
    cudaStream_t st;
    StageTimer(ibp_handle *h_, int stage_, cudaStream_t st_) : h(h_), stage(stage_), st(st_) {
        if (h->profiling) cudaEventRecord(h->ev[2 * stage], st);
    }
    ~StageTimer() {
        if (h->profiling) {
            cudaEventRecord(h->ev[2 * stage + 1], st);
            h->ev_used[stage] = 1;
        }
    }
};

}  // namespace

extern "C" {

int ibp_set_profiling(ibp_handle *h, int on) {
    if (!h) return fail(IBP_ERR_ARG, "ibp_set_profiling: null handle");
    CU(cudaSetDevice(h->device));
    if (on && !h->ev[0])
        for (int i = 0; i < 2 * IBP_NUM_STAGES; i++) CU(cudaEventCreate(&h->ev[i]));
    h->profiling = on ? 1 : 0;
    for (int i = 0; i < IBP_NUM_STAGES; i++) h->ev_used[i] = 0;
    return IBP_OK;
}

int ibp_get_stage_ms(ibp_handle *h, float *ms) {
    if (!h || !ms) return fail(IBP_ERR_ARG, "ibp_get_stage_ms: null argument");
    CU(cudaSetDevice(h->device));
    for (int i = 0; i < IBP_NUM_STAGES; i++) {
        ms[i] = -1.0f;
        if (h->profiling && h->ev_used[i]) {
            CU(cudaEventSynchronize(h->ev[2 * i + 1]));
            CU(cudaEventElapsedTime(&ms[i], h->ev[2 * i], h->ev[2 * i + 1]));
        }
    }
    return IBP_OK;
}

void ibp_default_params(ibp_params *p) {
    if (!p) return;
    p->limb_sample_thresh = 0.1f;
    p->step_ratio = 0.8f;
    p->min_parts = 2;
    p->min_human_score = 0.45f;
    p->max_steps = 20;
    p->len_rate = 16;
    p->merge_tolerance = 0.7f;
    p->weight_limb = 0.5f;
    p->weight_a = 0.25f;
    p->weight_b = 0.25f;
    p->peak_thresh = 0.1f;
    p->upsample = 4;
    p->refine_window = 2;
}

void ibp_default_limits(ibp_limits *l) {
    if (!l) return;
    l->max_batch = 64;
    l->max_peaks_per_part = 128;
    l->max_candidates_per_limb = 1024;
    l->max_bodies = 128;  // 128 * 160 B of grouping tables fit behind a 128x128 plane in the streaming limb kernel
    l->max_humans = 128;
    l->max_feat_h = 128;
    l->max_feat_w = 128;
}

const char *ibp_last_error(void) { return g_err.c_str(); }

int ibp_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

void ibp_destroy(ibp_handle *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    destroy_buffers(h);
    delete h;
}

int ibp_create(ibp_handle **out, int device, const ibp_params *params, const ibp_limits *limits) {
    if (!out) return fail(IBP_ERR_ARG, "ibp_create: out is NULL");
    *out = nullptr;
    int ndev = ibp_device_count();
    if (ndev <= 0) return fail(IBP_ERR_NO_DEVICE, "no CUDA device visible: libibp_b200 has no CPU path");
    if (device < 0 || device >= ndev) return fail(IBP_ERR_ARG, "device %d out of range [0,%d)", device, ndev);
    ibp_handle *h = new ibp_handle();
    h->device = device;
    if (params) h->P = *params; else ibp_default_params(&h->P);
    if (limits) h->L = *limits; else ibp_default_limits(&h->L);
    int rc = validate_limits(h->L);
    if (rc != IBP_OK) {
        delete h;
        return rc;
    }
    if (h->P.upsample < 1 || h->P.refine_window < 1 || h->P.refine_window > 8 || h->P.max_steps < 2) {
        delete h;
        return fail(IBP_ERR_ARG, "ibp_params out of range");
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        delete h;
        return fail(IBP_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    }
    const size_t B = h->L.max_batch, cap = h->L.max_peaks_per_part;
    h->line_stride = (int)(cap * kParts);
    auto alloc = [&]() -> int {
        CU(cudaMalloc(&h->d_peaks, B * kParts * cap * sizeof(DevPeak)));
        CU(cudaMalloc(&h->d_subpix, B * kParts * cap * 2 * sizeof(double)));
        CU(cudaMalloc(&h->d_part_count, B * kParts * sizeof(int)));
        CU(cudaMalloc(&h->d_links, B * kLimbs * cap * sizeof(DevLink)));
        CU(cudaMalloc(&h->d_link_count, B * kLimbs * sizeof(int)));
        CU(cudaMalloc(&h->d_line, B * h->line_stride * sizeof(ibp_peak)));
        CU(cudaMalloc(&h->d_n_peaks, B * sizeof(int)));
        CU(cudaMalloc(&h->d_humans, B * h->L.max_humans * sizeof(ibp_human)));
        CU(cudaMalloc(&h->d_n_humans, B * sizeof(int)));
        // status words, then the per-image raw peak counts, then the streaming limb kernel's image tickets: one region,
        // zeroed by ONE memset at the start of every pipeline launch (a ticket left behind by a stalled launch cannot
        // reach the next one)
        CU(cudaMalloc(&h->d_status, 3 * B * sizeof(uint32_t)));
        h->d_n_raw = reinterpret_cast<int *>(h->d_status + B);
        CU(cudaMemset(h->d_part_count, 0, B * kParts * sizeof(int)));
        CU(cudaMemset(h->d_link_count, 0, B * kLimbs * sizeof(int)));
        CU(cudaMemset(h->d_n_peaks, 0, B * sizeof(int)));
        CU(cudaMemset(h->d_n_humans, 0, B * sizeof(int)));
        CU(cudaMemset(h->d_status, 0, 3 * B * sizeof(uint32_t)));
        CU(cudaMalloc(&h->d_ticket, (2 * B + 1) * sizeof(unsigned)));
        CU(cudaMemset(h->d_ticket, 0, (2 * B + 1) * sizeof(unsigned)));
        CU(cudaMalloc(&h->d_order, B * sizeof(int)));
        CU(cudaMalloc(&h->d_partbase, B * (kParts + 1) * sizeof(int)));
        CU(cudaMalloc(&h->d_item_meta, B * (kLimbs + 1) * sizeof(int4)));
        CU(cudaMalloc(&h->d_yext, B * kParts * 4 * sizeof(int)));
        CU(cudaMemset(h->d_yext, 0x80, B * kParts * 4 * sizeof(int)));  // 0x80808080: below every -y
        CU(cudaMalloc(&h->d_lb_work, sizeof(unsigned)));
        CU(cudaMemset(h->d_lb_work, 0, sizeof(unsigned)));
        CU(cudaFuncSetAttribute(limbs_stream_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_stream_kernel<__half, false, kLimbStreamThreadsHalf>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_stream_kernel<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_stream_kernel<__half, true, kLimbStreamThreadsHalf>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaMalloc(&h->d_pk_work, sizeof(unsigned)));
        CU(cudaMemset(h->d_pk_work, 0, sizeof(unsigned)));
        CU(cudaMallocHost(&h->h_n_peaks, B * sizeof(int)));
        CU(cudaMallocHost(&h->h_part_count, B * kParts * sizeof(int)));
        CU(cudaMallocHost(&h->h_line, B * h->line_stride * sizeof(ibp_peak)));
        CU(cudaMallocHost(&h->h_n_humans, B * sizeof(int)));
        CU(cudaMallocHost(&h->h_humans, B * h->L.max_humans * sizeof(ibp_human)));
        CU(cudaMallocHost(&h->h_status, B * sizeof(uint32_t)));
        CU(cudaMallocHost(&h->h_peaks_stage, kParts * cap * sizeof(DevPeak)));
        memset(h->h_n_peaks, 0, B * sizeof(int));
        memset(h->h_n_humans, 0, B * sizeof(int));
        memset(h->h_status, 0, B * sizeof(uint32_t));
        // opt in to large dynamic shared memory once
        CU(cudaFuncSetAttribute(peaks_kernel<float, IBP_PEAKS_CROSS_REFINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_kernel<__half, IBP_PEAKS_CROSS_REFINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_kernel<float, IBP_PEAKS_BOX_CENTROID>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_kernel<__half, IBP_PEAKS_BOX_CENTROID>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, true, __half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, false, __half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_stream_kernel<IBP_PEAKS_BOX_CENTROID, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(peaks_stream_kernel<IBP_PEAKS_BOX_CENTROID, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, h->device));
        CU(cudaFuncSetAttribute(limbs_stream3_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_stream3_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        if (h->L.max_peaks_per_part <= 128 && !getenv("IBP_LIMBS_NO_STREAM3"))  // the handles limbs_stream3_kernel serves
            CU(cudaMalloc(&h->d_l3_scratch, (size_t)kL3MaxCtasPerSm * h->sm_count * l3_global_scratch_bytes(h->L.max_candidates_per_limb)));
        CU(cudaFuncSetAttribute(limbs_kernel<float, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_kernel<__half, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_kernel<float, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(limbs_kernel<__half, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(upsample_limbs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(fuse_multiscale_kernel<kMsStride, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(fuse_multiscale_kernel<kMsStride, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(fuse_multiscale_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        CU(cudaFuncSetAttribute(fuse_multiscale_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
        if (limb_phase_bytes(h, 0) > kMaxDynSmem)
            return fail(IBP_ERR_ARG, "limits need %zu B of shared memory for the limb kernel (max %zu)", limb_phase_bytes(h, 0), kMaxDynSmem);
        if (group_phase_bytes(h) > kMaxDynSmem)
            return fail(IBP_ERR_ARG, "max_bodies=%d needs %zu B of shared memory (max %zu)", h->L.max_bodies, group_phase_bytes(h), kMaxDynSmem);
        // the fills above ran on the legacy default stream, which a cudaStreamNonBlocking stream (torch's) never waits
        // for: they are complete before the first launch on any user stream
        CU(cudaDeviceSynchronize());
        return IBP_OK;
    };
    rc = alloc();
    if (rc != IBP_OK) {
        std::string keep = g_err;
        destroy_buffers(h);
        delete h;
        g_err = keep;
        return rc;
    }
    *out = h;
    return IBP_OK;
}

int64_t ibp_launch_count(const ibp_handle *h) { return h ? h->launches : 0; }

// ------------------------------------------------------------------------------------------------ K1
int ibp_fuse_flip(ibp_handle *h, const void *net_out, int dtype, int batch, int fh, int fw, int flip_avg, float *fused, void *stream) {
    if (!h || !net_out || !fused) return fail(IBP_ERR_ARG, "ibp_fuse_flip: null argument");
    if (batch < 1 || fh < 1 || fw < 1) return fail(IBP_ERR_SHAPE, "ibp_fuse_flip: bad shape (%d,%d,%d)", batch, fh, fw);
    if (dtype != IBP_F32 && dtype != IBP_F16) return fail(IBP_ERR_ARG, "ibp_fuse_flip: dtype %d", dtype);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t hw = (size_t)fh * fw;
    const bool aligned = (((uintptr_t)net_out) & 15) == 0 && (((uintptr_t)fused) & 15) == 0;
    if (dtype == IBP_F32 && aligned && fw % 4 == 0) {
        const int hw4 = (int)(hw / 4);
        dim3 grid((hw4 + 255) / 256, kLayers, batch);
        StageTimer t(h, IBP_STAGE_FUSE, st);
        fuse_flip_f32_vec4<<<grid, 256, 0, st>>>((const float *)net_out, fused, hw4, fw / 4, flip_avg);
    } else if (dtype == IBP_F16 && aligned && fw % 8 == 0) {
        const int hw8 = (int)(hw / 8);
        dim3 grid((hw8 + 255) / 256, kLayers, batch);
        StageTimer t(h, IBP_STAGE_FUSE, st);
        fuse_flip_f16_vec8<<<grid, 256, 0, st>>>((const __half *)net_out, fused, hw8, fw / 8, flip_avg);
    } else {
        dim3 grid((unsigned)std::min<size_t>((hw + 255) / 256, 65535), kLayers, batch);
        StageTimer t(h, IBP_STAGE_FUSE, st);
        if (dtype == IBP_F32)
            fuse_flip_scalar<float><<<grid, 256, 0, st>>>((const float *)net_out, fused, fh, fw, flip_avg);
        else
            fuse_flip_scalar<__half><<<grid, 256, 0, st>>>((const __half *)net_out, fused, fh, fw, flip_avg);
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}


// ------------------------------------------------------------------------------------------------ K1 multi-scale, strip form
namespace {

// Host mirror of cubic_weights / ms_tap (ibp_device.cuh, k_multiscale.cuh): plain IEEE float and double operations in the
// same order, no contraction (the library is built with -ffp-contract=off for the host too), so the tables hold the
// very bits the device would compute.
void host_cubic_weights(float t, float *w) {
    const float A = -0.75f;
    const float t1 = t + 1.0f, u = 1.0f - t;
    float x = A * t1;
    x = x - 5.0f * A;
    x = x * t1;
    x = x + 8.0f * A;
    x = x * t1;
    w[0] = x - 4.0f * A;
    x = (A + 2.0f) * t;
    x = x - (A + 3.0f);
    x = x * t;
    x = x * t;
    w[1] = x + 1.0f;
    x = (A + 2.0f) * u;
    x = x - (A + 3.0f);
    x = x * u;
    x = x * u;
    w[2] = x + 1.0f;
    x = 1.0f - w[0];
    x = x - w[1];
    w[3] = x - w[2];
}
int host_ms_tap(int d, double scale, float *w) {
    double t = (double)d + 0.5;
    t = t * scale;
    t = t - 0.5;
    float f = (float)t;
    const int s = (int)floorf(f);
    f = f - (float)s;
    host_cubic_weights(f, w);
    return s;
}
inline int floor_div4(int v) { return v >= 0 ? v / 4 : -((-v + 3) / 4); }

// Returns IBP_OK with *used = true when the strip kernel took the call; *used = false when the shape is outside what it
// handles (the tile kernel then runs).
int launch_fuse_strip(ibp_handle *h, const MsArgs &m, int batch, cudaStream_t st, bool *used) {
    *used = false;
    static const char *force = getenv("IBP_MS_KERNEL");  // "tile" / "strip": experiments and the tests that hold both to the goldens
    if (!force || strcmp(force, "strip")) return IBP_OK;  // opt-in until it beats the tile kernel
    const int H = m.H, W = m.W, ns = m.n_scales;
    if (H < kSsRows || (size_t)batch * kLayers > 65535u) return IBP_OK;
    const int n_strips = (W + kSsCols - 1) / kSsCols;

    // per-row tables of the whole image, and from them the exact sizes of the rolling buffers
    std::vector<SsYEntry> ytab((size_t)ns * H);
    std::vector<SsXEntry> xtab((size_t)ns * W);
    auto sa = std::make_unique<SsArgs>();
    SsArgs &a = *sa;
    memset(&a, 0, sizeof(SsArgs) - sizeof(a.ytab));
    int nc_max = 1, cap_max = 2, s1_max = 4, nfx_max = 4, hist_floats = 0;
    for (int s = 0; s < ns; s++) {
        const MsScale &S = m.sc[s];
        for (int Y = 0; Y < H; Y++) {
            SsYEntry &e = ytab[(size_t)s * H + Y];
            const int y0 = host_ms_tap(Y, S.scale_y, e.w) - 1;
            e.g = floor_div4(y0 + 1);
            e.o = y0 - (4 * e.g - 2);
        }
        for (int X = 0; X < W; X++) {
            SsXEntry &e = xtab[(size_t)s * W + X];
            e.xv0 = host_ms_tap(X, S.scale_x, e.w) - 1;
            e.pad[0] = e.pad[1] = e.pad[2] = 0;
        }
        int n_kx_s = 1;
        for (int X0 = 0; X0 < W; X0 += kSsCols) {
            const int x_lo = xtab[(size_t)s * W + X0].xv0, x_hi = xtab[(size_t)s * W + std::min(X0 + kSsCols - 1, W - 1)].xv0 + 3;
            nc_max = std::max(nc_max, x_hi - x_lo + 1);
            const int kx_lo = floor_div4(std::min(std::max(x_lo, 0), S.sw - 1) - 2), kx_hi = floor_div4(std::min(std::max(x_hi, 0), S.sw - 1) - 2);
            n_kx_s = std::max(n_kx_s, kx_hi - kx_lo + 1);
        }
        SsScale &D = a.sc[s];
        D.src = S.src;
        D.h = S.h, D.w = S.w, D.sh = S.sh, D.sw = S.sw;
        D.s1 = 4 * n_kx_s;
        D.hist_off = hist_floats;
        D.kprime = floor_div4(S.sh - 3);
        hist_floats += kSsHist * D.s1;
        s1_max = std::max(s1_max, D.s1);
        nfx_max = std::max(nfx_max, n_kx_s + 3);
    }
    // rows per launch (the per-row table travels in the parameter space) and per warp
    int rows_cap = (kSsYCap / ns) & ~(kSsRows - 1);
    if (rows_cap < kSsRows) return IBP_OK;
    const long long items_wanted = 16LL * h->sm_count * 16;  // ~16 waves of resident warps
    const long long per_chain = (long long)n_strips * kLayers * batch;
    long long chains = std::max(1LL, (items_wanted + per_chain - 1) / per_chain);
    int chain_rows = (int)((H + chains - 1) / chains);
    chain_rows = std::max(40, chain_rows);
    chain_rows = std::min((chain_rows + kSsRows - 1) & ~(kSsRows - 1), rows_cap);
    const int chunk_rows = (rows_cap / chain_rows) * chain_rows;
    // row groups a phase produces: exact, from the table, for the segment boundaries this launch plan creates
    for (int s = 0; s < ns; s++) {
        const SsYEntry *yt = &ytab[(size_t)s * H];
        int cap = 2;
        for (int c0 = 0; c0 < H; c0 += chunk_rows)
            for (int Yc = c0; Yc < std::min(c0 + chunk_rows, H); Yc += chain_rows) {
                const int Yc1 = std::min(std::min(Yc + chain_rows, c0 + chunk_rows), H);
                int g_cur = yt[Yc].g;
                for (int Ys = Yc; Ys < Yc1; Ys += kSsRows) {
                    const int g_need = yt[std::min(Ys + kSsRows, Yc1) - 1].g;
                    cap = std::max(cap, g_need - g_cur);
                    g_cur = g_need;
                }
            }
        a.sc[s].cap = cap;
        cap_max = std::max(cap_max, cap);
    }
    if (cap_max > kSsMaxNew) return IBP_OK;  // a down-scale this steep: the tile kernel
    const int t1_rows = kSsHist + std::max(cap_max, 5);
    const int slots = cap_max | 1;  // an odd number of 16-byte slots per column: conflict-free 128-bit stores
    const int it_stride = 16 * slots;
    a.f_stride = (nfx_max + 1) & ~1;
    // I_T first; stage A's feature rows share its space (B has consumed them before C writes I_T)
    size_t off = std::max((size_t)nc_max * it_stride, (size_t)kSsMaxNew * a.f_stride * sizeof(float));
    off = (off + 15) & ~(size_t)15;
    a.off_t1 = (int)off;
    off += (size_t)t1_rows * s1_max * sizeof(float);
    a.off_hist = (int)off;
    off += (size_t)hist_floats * sizeof(float);
    a.off_tw = (int)off;
    off += 32 * sizeof(float) + 2 * kMsMaxScales * sizeof(int) + kMsMaxScales * sizeof(int4);
    a.off_pf = (int)off;
    off += 2 * (size_t)kSsMaxNew * a.f_stride * sizeof(float);
    const size_t smem = off;
    if (smem > 48 * 1024) return IBP_OK;

    // per-column tables on the device (cached by shape)
    std::vector<int> key = {W, ns};
    for (int s = 0; s < ns; s++) key.push_back(m.sc[s].sw);
    const size_t n_x = (size_t)ns * W;
    if (key != h->ms_xtab_key) {
        if (h->ms_xtab_cap < n_x) {
            CU(cudaDeviceSynchronize());
            free_dev(h->d_ms_xtab);
            free_host(h->h_ms_xtab);
            h->ms_xtab_cap = 0;
            CU(cudaMalloc(&h->d_ms_xtab, n_x * sizeof(SsXEntry)));
            CU(cudaMallocHost(&h->h_ms_xtab, n_x * sizeof(SsXEntry)));
            h->ms_xtab_cap = n_x;
        } else {
            CU(cudaStreamSynchronize(st));  // the staging buffer may still feed an earlier copy
        }
        memcpy(h->h_ms_xtab, xtab.data(), n_x * sizeof(SsXEntry));
        CU(cudaMemcpyAsync(h->d_ms_xtab, h->h_ms_xtab, n_x * sizeof(SsXEntry), cudaMemcpyHostToDevice, st));
        h->ms_xtab_key = key;
    }
    for (int s = 0; s < ns; s++) a.sc[s].xtab = h->d_ms_xtab + (size_t)s * W;

    a.n_scales = ns;
    a.flip = m.flip;
    a.H = H;
    a.W = W;
    a.one = 1.0f;
    a.out = m.out;
    a.heat32 = m.heat32;
    a.chain_rows = chain_rows;
    const bool pow2 = (ns & (ns - 1)) == 0;
    // rows of the interleaved 30- / 20-channel images that do not end on a multiple of four elements have a scalar tail
    const bool tail = ((W * kLimbs) & 3) != 0 || ((W * (kLayers - kLimbs)) & 3) != 0;
    {
        StageTimer t(h, IBP_STAGE_FUSE, st);
        for (int y0 = 0; y0 < H; y0 += chunk_rows) {
            const int y1 = std::min(y0 + chunk_rows, H), n = y1 - y0;
            a.y0 = y0;
            a.y1 = y1;
            for (int s = 0; s < ns; s++) memcpy(&a.ytab[(size_t)s * n], &ytab[(size_t)s * H + y0], (size_t)n * sizeof(SsYEntry));
            const dim3 grid(n_strips, (n + chain_rows - 1) / chain_rows, (unsigned)batch * kLayers);
            void (*kern)(const SsArgs) = nullptr;
#define IBP_SS_PICK(N) (tail ? (pow2 ? fuse_strip_kernel<N, true, true> : fuse_strip_kernel<N, false, true>) \
                             : (pow2 ? fuse_strip_kernel<N, true, false> : fuse_strip_kernel<N, false, false>))
            switch (slots) {
                case 3: kern = IBP_SS_PICK(3); break;
                case 5: kern = IBP_SS_PICK(5); break;
                case 7: kern = IBP_SS_PICK(7); break;
                default: kern = IBP_SS_PICK(9); break;
            }
#undef IBP_SS_PICK
            kern<<<grid, kSsCols, smem, st>>>(a);
            h->launches++;
        }
    }
    CU(cudaGetLastError());
    *used = true;
    return IBP_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ K1 multi-scale
int ibp_fuse_multiscale(ibp_handle *h, const ibp_scale_input *scales, int n_scales, int batch, int img_h, int img_w, int flip_avg,
                        double *out, float *heat32, void *stream) {
    if (!h || !scales || !out) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale: null argument");
    if (n_scales < 1 || n_scales > kMsMaxScales) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale: 1..%d scales, got %d", kMsMaxScales, n_scales);
    if (batch < 1 || img_h < 1 || img_w < 1) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale: bad shape");
    if (h->P.upsample != 4) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale: supports stride 4 only (params.upsample = %d)", h->P.upsample);
    CU(cudaSetDevice(h->device));
    MsArgs a;
    memset(&a, 0, sizeof(a));
    a.n_scales = n_scales;
    a.up = h->P.upsample;
    a.flip = flip_avg ? 1 : 0;
    a.one = 1.0f;
    a.H = img_h;
    a.W = img_w;
    a.out = out;
    a.heat32 = heat32;
    size_t cap_r1 = 4, cap_i = 4, max_nkx = 1;
    for (int s = 0; s < n_scales; s++) {
        const ibp_scale_input &in = scales[s];
        if (!in.net_out || in.fh < 1 || in.fw < 1 || in.sh < 1 || in.sw < 1 || in.sh > in.fh * a.up || in.sw > in.fw * a.up)
            return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale: scale %d: feature (%d,%d), scaled image (%d,%d)", s, in.fh, in.fw, in.sh, in.sw);
        MsScale &S = a.sc[s];
        S.src = in.net_out;
        S.h = in.fh;
        S.w = in.fw;
        S.sh = in.sh;
        S.sw = in.sw;
        // cv2.resize with an explicit dsize: inv_scale = (double)dsize / ssize; scale = 1. / inv_scale
        S.scale_x = 1.0 / ((double)img_w / (double)in.sw);
        S.scale_y = 1.0 / ((double)img_h / (double)in.sh);
        max_nkx = std::max(max_nkx, ms_footprint(S.scale_x, S.scale_y, 0).n_kx);
    }
    {
        bool used = false;
        const int rc = launch_fuse_strip(h, a, batch, (cudaStream_t)stream, &used);
        if (rc != IBP_OK || used) return rc;
    }
    // every scale narrow enough: the variant with compile-time row strides
    const size_t stride = (4 * max_nkx <= (size_t)kMsStride) ? kMsStride : 0;
    for (int s = 0; s < n_scales; s++) {
        const MsFootprint f = ms_footprint(a.sc[s].scale_x, a.sc[s].scale_y, stride);
        cap_r1 = std::max(cap_r1, f.r1);
        cap_i = std::max(cap_i, f.i);
    }
    a.cap_r1 = (int)cap_r1;
    a.cap_i = (int)cap_i;
    const size_t smem = (cap_r1 + cap_i) * sizeof(float);
    if (smem > kMaxDynSmem) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale: a scale factor this large needs %zu B of shared memory", smem);
    const dim3 grid((img_w + kMsTileX - 1) / kMsTileX, (img_h + kMsTileY - 1) / kMsTileY, (unsigned)batch * kLayers);
    if (grid.z > 65535u) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale: batch %d too large for one launch (max %d)", batch, 65535 / kLayers);
    {
        StageTimer t(h, IBP_STAGE_FUSE, (cudaStream_t)stream);
        const bool pow2 = (n_scales & (n_scales - 1)) == 0;
        auto kern = stride ? (pow2 ? fuse_multiscale_kernel<kMsStride, true> : fuse_multiscale_kernel<kMsStride, false>)
                           : (pow2 ? fuse_multiscale_kernel<0, true> : fuse_multiscale_kernel<0, false>);
        kern<<<grid, kMsThreads, smem, (cudaStream_t)stream>>>(a);
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}


// ------------------------------------------------------------------------------------------------ second operator mode
namespace {

struct Lin1 {
    int i0, i1;
    float a0, a1;
};
// cv2.resize INTER_LINEAR, one axis: source coordinate (d + 0.5) * scale - 0.5 in float, taps floor and floor + 1, the
// fraction forced to 0 where the pair would leave the source (resize.cpp, the xofs / alpha set-up of the generic path)
Lin1 cv_linear_tap(int d, double scale, int n_src) {
    double t = (double)d + 0.5;
    t = t * scale;
    t = t - 0.5;
    float f = (float)t;
    int s = (int)floorf(f);
    f = f - (float)s;
    if (s < 0) f = 0.f, s = 0;
    if (s >= n_src - 1) f = 0.f, s = n_src - 1;
    return {s, std::min(s + 1, n_src - 1), 1.0f - f, f};
}
// destination index d of the second resize (crop of n_crop -> n_dst) composed with the x4 resize (n_feat -> 4 n_feat)
bool lin_compose(int d, double scale2, int n_crop, int n_feat, float fold, LinTap *out) {
    const Lin1 t2 = cv_linear_tap(d, scale2, n_crop);
    int idx[4];
    float wgt[4];
    const int u[2] = {t2.i0, t2.i1};
    const float b[2] = {t2.a0, t2.a1};
    for (int k = 0; k < 2; k++) {
        const Lin1 t1 = cv_linear_tap(u[k], 0.25, n_feat);
        idx[2 * k] = t1.i0, wgt[2 * k] = b[k] * t1.a0;
        idx[2 * k + 1] = t1.i1, wgt[2 * k + 1] = b[k] * t1.a1;
    }
    int first = idx[0];
    for (int k = 1; k < 4; k++) first = std::min(first, idx[k]);
    const int base = std::max(0, std::min(first, n_feat - 3));
    out->base = base;
    out->c[0] = out->c[1] = out->c[2] = 0.f;
    for (int k = 0; k < 4; k++) {
        const int j = idx[k] - base;
        if (j < 0 || j > 2) return false;
        out->c[j] += wgt[k] * fold;
    }
    return true;
}

}  // namespace

int ibp_fuse_multiscale_linear(ibp_handle *h, const ibp_scale_input *scales, int n_scales, int batch, int img_h, int img_w, int flip_avg,
                               double *out, float *heat32, void *stream) {
    if (!h || !scales || !out) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale_linear: null argument");
    if (n_scales < 1 || n_scales > kLinMaxScales) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale_linear: 1..%d scales, got %d", kLinMaxScales, n_scales);
    if (batch < 1 || img_h < 1 || img_w < 1 || (size_t)batch * kLayers > 65535u) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: bad shape");
    if (h->P.upsample != 4) return fail(IBP_ERR_ARG, "ibp_fuse_multiscale_linear: supports stride 4 only (params.upsample = %d)", h->P.upsample);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int H = img_h, W = img_w, ns = n_scales;
    std::vector<int> key = {H, W, ns};
    for (int s = 0; s < ns; s++) {
        const ibp_scale_input &in = scales[s];
        if (!in.net_out || in.fh < 3 || in.fw < 4 || in.sh < 1 || in.sw < 1 || in.sh > in.fh * 4 || in.sw > in.fw * 4)
            return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: scale %d: feature (%d,%d), scaled image (%d,%d)", s, in.fh, in.fw, in.sh, in.sw);
        if ((in.fw & 3) || (((uintptr_t)in.net_out) & 15))  // the rows are staged in 16-byte pieces; every map the reference pads to x64 qualifies
            return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: scale %d: feature width %d must be a multiple of 4 and the map 16-byte aligned", s, in.fw);
        key.insert(key.end(), {in.fh, in.fw, in.sh, in.sw});
    }
    const int n_seg_all = (H + kLinSeg - 1) / kLinSeg;
    const size_t per_scale = (size_t)W + H + (n_seg_all + 3) / 4;  // taps of the columns, of the rows, then the segments' last feature rows (ints)
    const size_t n_tab = (size_t)ns * per_scale;
    if (key != h->lin_tab_key) {
        std::vector<LinTap> tab(n_tab);
        const float fold = 1.0f / (float)ns;  // "/ n_scales" of parse_skeletons.py:280-281, folded into the row weights
        for (int s = 0; s < ns; s++) {
            const ibp_scale_input &in = scales[s];
            const double sx = 1.0 / ((double)W / (double)in.sw), sy = 1.0 / ((double)H / (double)in.sh);
            LinTap *xt = &tab[(size_t)s * per_scale], *yt = xt + W;
            for (int X = 0; X < W; X++)
                if (!lin_compose(X, sx, in.sw, in.fw, 1.0f, &xt[X])) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: taps of column %d span more than three", X);
            for (int Y = 0; Y < H; Y++)
                if (!lin_compose(Y, sy, in.sh, in.fh, fold, &yt[Y])) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: taps of row %d span more than three", Y);
            int *seg_hi = reinterpret_cast<int *>(yt + H);
            for (int g = 0; g < n_seg_all; g++) seg_hi[g] = yt[std::min((g + 1) * kLinSeg, H) - 1].base + 2;
        }
        if (h->lin_tab_cap < n_tab) {
            CU(cudaDeviceSynchronize());
            free_dev(h->d_lin_tab);
            free_host(h->h_lin_tab);
            h->lin_tab_cap = 0;
            CU(cudaMalloc(&h->d_lin_tab, n_tab * sizeof(LinTap)));
            CU(cudaMallocHost(&h->h_lin_tab, n_tab * sizeof(LinTap)));
            h->lin_tab_cap = n_tab;
        } else {
            CU(cudaStreamSynchronize(st));  // the staging buffer may still feed an earlier copy
        }
        memcpy(h->h_lin_tab, tab.data(), n_tab * sizeof(LinTap));
        CU(cudaMemcpyAsync(h->d_lin_tab, h->h_lin_tab, n_tab * sizeof(LinTap), cudaMemcpyHostToDevice, st));
        h->lin_tab_key = key;
    }
    LinArgs a;
    memset(&a, 0, sizeof(a));
    for (int s = 0; s < ns; s++) {
        a.sc[s].src = scales[s].net_out;
        a.sc[s].h = scales[s].fh;
        a.sc[s].w = scales[s].fw;
        a.sc[s].xtab = h->d_lin_tab + (size_t)s * per_scale;
        a.sc[s].ytab = a.sc[s].xtab + W;
        a.sc[s].seg_hi = reinterpret_cast<const int *>(a.sc[s].ytab + H);
    }
    a.n_scales = ns;
    a.flip = flip_avg ? 1 : 0;
    a.H = H;
    a.W = W;
    a.out = out;
    a.heat32 = heat32;
    // rows per CTA: enough chains for ~8 waves of 8 resident CTAs per SM, whole segments, at least 32 rows each (a chain start
    // costs three feature rows per scale)
    // 256 or 128 output columns per CTA: what wastes fewer columns right of the image (640 = 5 x 128)
    const int lin_threads = (((W + 255) / 256) * 256 - W > ((W + 127) / 128) * 128 - W) ? 64 : 128;
    const int kLinCols = 2 * lin_threads;
    const long long cols = (W + kLinCols - 1) / kLinCols;
    const long long per_chain = cols * kLayers * batch;
    const long long want = 8LL * 8 * h->sm_count;
    const long long chains = std::max(1LL, std::min((long long)(H + 31) / 32, (want + per_chain - 1) / per_chain));
    a.chain_rows = (int)((H + chains - 1) / chains);
    a.chain_rows = std::min((a.chain_rows + kLinSeg - 1) / kLinSeg * kLinSeg, kLinSeg * kLinMaxSegs);
    // exact capacities of the staging buffers for this launch plan, from the (cached) host copy of the tables
    {
        const LinTap *tab = h->h_lin_tab;
        int off = 0, hoff = 0;
        a.step1 = a.ystep1 = 1;
        for (int s = 0; s < ns; s++) {
            const LinTap *xt = tab + (size_t)s * per_scale, *yt = xt + W;
            int nc = 1, maxr = 1;
            // staged columns of a CTA: from its first tap rounded down to a 16-byte boundary, whole 16-byte groups
            for (int X0 = 0; X0 < W; X0 += kLinCols) nc = std::max(nc, ((xt[std::min(X0 + kLinCols - 1, W - 1)].base + 2 - (xt[X0].base & ~3) + 1) + 3) & ~3);
            if (nc / 4 > lin_threads) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: scale %d shrinks by more than a factor of ~7.9", s);
            for (int Yc = 0; Yc < H; Yc += a.chain_rows) {
                int loaded = yt[Yc].base - 1;
                for (int Ya = Yc; Ya < std::min(Yc + a.chain_rows, H); Ya += kLinSeg) {
                    const int f_hi = yt[std::min(std::min(Ya + kLinSeg, Yc + a.chain_rows), H) - 1].base + 2;
                    maxr = std::max(maxr, f_hi - loaded);
                    loaded = std::max(loaded, f_hi);
                }
            }
            for (int X = 0; X + 1 < W; X++)
                if (xt[X + 1].base - xt[X].base > 1) a.step1 = 0;
            for (int Y = 0; Y + 1 < H; Y++)
                if (yt[Y + 1].base - yt[Y].base > 1) a.ystep1 = 0;
            nc += 4;  // the four-load form of the horizontal pass reads one column past a thread's last tap
            a.nc[s] = nc;
            a.maxr[s] = maxr;
            a.off[s] = off;
            off += 2 * nc * maxr;
            a.hoff[s] = hoff;
            hoff += (maxr + 4) * kLinCols;  // three rows kept from the segment before, one row of slack for the paired vertical pass
        }
        a.buf_floats = (off + 3) & ~3;
        a.hr_floats = hoff;
    }
    const size_t smem = ((size_t)a.buf_floats + a.hr_floats) * sizeof(float) + 2 * (size_t)ns * kLinSeg * sizeof(LinTap);
    if (smem > 160 * 1024) return fail(IBP_ERR_SHAPE, "ibp_fuse_multiscale_linear: staging buffers of %zu B", smem);
    const dim3 grid((unsigned)cols, (unsigned)((H + a.chain_rows - 1) / a.chain_rows), (unsigned)batch * kLayers);
#define IBP_LIN_LAUNCH(N)                                                                                          \
    do {                                                                                                           \
        if (smem > 48 * 1024) {                                                                                    \
            CU(cudaFuncSetAttribute(fuse_linear_kernel<N, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));  \
            CU(cudaFuncSetAttribute(fuse_linear_kernel<N, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)); \
        }                                                                                                          \
        if (lin_threads == 64) fuse_linear_kernel<N, 64><<<grid, 64, smem, st>>>(a);                               \
        else fuse_linear_kernel<N, 128><<<grid, 128, smem, st>>>(a);                                               \
    } while (0)
    {
        StageTimer t(h, IBP_STAGE_FUSE, st);
        switch (ns) {
            case 1: IBP_LIN_LAUNCH(1); break;
            case 2: IBP_LIN_LAUNCH(2); break;
            case 3: IBP_LIN_LAUNCH(3); break;
            case 4: IBP_LIN_LAUNCH(4); break;
            case 5: IBP_LIN_LAUNCH(5); break;
            case 6: IBP_LIN_LAUNCH(6); break;
            case 7: IBP_LIN_LAUNCH(7); break;
            default: IBP_LIN_LAUNCH(8); break;
        }
    }
#undef IBP_LIN_LAUNCH
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

int ibp_smooth_gaussian(ibp_handle *h, const float *maps, float *out, int planes, int img_h, int img_w, int ksize, float sigma, void *stream) {
    if (!h || !maps || !out) return fail(IBP_ERR_ARG, "ibp_smooth_gaussian: null argument");
    if (planes < 1 || planes > 65535 || img_h < 1 || img_w < 1) return fail(IBP_ERR_SHAPE, "ibp_smooth_gaussian: bad shape");
    if (ksize < 1 || ksize > kGsMaxK || !(ksize & 1) || !(sigma > 0.f)) return fail(IBP_ERR_ARG, "ibp_smooth_gaussian: odd kernel size 1..%d and sigma > 0", kGsMaxK);
    if (img_h <= ksize / 2 || img_w <= ksize / 2) return fail(IBP_ERR_SHAPE, "ibp_smooth_gaussian: the mirrored border needs a map larger than half the kernel");
    CU(cudaSetDevice(h->device));
    GaussArgs a;
    memset(&a, 0, sizeof(a));
    a.src = maps;
    a.dst = out;
    a.H = img_h;
    a.W = img_w;
    a.ksize = ksize;
    // utils/util.py:131-146 in float32 like torch: g[i] = 1 / (sigma sqrt(2 pi)) exp(-((i - mean) / sigma)^2 / 2), the 2-D kernel is the
    // outer product divided by its sum = the product of the two normalised 1-D kernels
    {
        const float mean = (float)(ksize - 1) / 2.0f;
        float g[kGsMaxK], sum = 0.f;
        for (int i = 0; i < ksize; i++) {
            const float z = ((float)i - mean) / sigma;
            g[i] = 1.0f / (sigma * sqrtf(2.0f * 3.14159265358979323846f)) * expf(-(z * z) / 2.0f);
            sum += g[i];
        }
        for (int i = 0; i < ksize; i++) a.g[i] = g[i] / sum;
    }
    const int p = ksize / 2;
    // the 128-bit variant: sizes 3..9, rows that are 16-byte multiples on a 16-byte aligned map
    const bool vec = ksize >= 3 && ksize <= 9 && (img_w & 3) == 0 && (((uintptr_t)maps | (uintptr_t)out) & 15) == 0;
    const size_t smem = vec ? (size_t)(kGsTileY + 2 * p) * kGsTileX * sizeof(float)
                            : ((size_t)(kGsTileY + 2 * p) * (kGsTileX + 2 * p) + (size_t)(kGsTileY + 2 * p) * kGsTileX) * sizeof(float);
    static std::once_flag once;
    std::call_once(once, [] {
        cudaFuncSetAttribute(gauss_smooth_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        cudaFuncSetAttribute(gauss_smooth_kernel<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    });
    const dim3 grid((img_w + kGsTileX - 1) / kGsTileX, (img_h + kGsTileY - 1) / kGsTileY, planes);
    {
        StageTimer t(h, IBP_STAGE_SMOOTH, (cudaStream_t)stream);
        switch (vec ? ksize : 0) {
            case 3: gauss_smooth_kernel<3><<<grid, 256, smem, (cudaStream_t)stream>>>(a); break;
            case 5: gauss_smooth_kernel<5><<<grid, 256, smem, (cudaStream_t)stream>>>(a); break;
            case 7: gauss_smooth_kernel<7><<<grid, 256, smem, (cudaStream_t)stream>>>(a); break;
            case 9: gauss_smooth_kernel<9><<<grid, 256, smem, (cudaStream_t)stream>>>(a); break;
            default: gauss_smooth_kernel<0><<<grid, 256, smem, (cudaStream_t)stream>>>(a); break;
        }
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

// ------------------------------------------------------------------------------------------------ A4
int ibp_upsample_limbs(ibp_handle *h, const float *fused, int chan_stride, int batch, int fh, int fw, int up, float *out_hwc, void *stream) {
    if (!h || !fused || !out_hwc) return fail(IBP_ERR_ARG, "ibp_upsample_limbs: null argument");
    if (batch < 1 || fh < 1 || fw < 1 || up < 1 || chan_stride < kLimbs) return fail(IBP_ERR_SHAPE, "ibp_upsample_limbs: bad shape");
    CU(cudaSetDevice(h->device));
    UpArgs a;
    a.src = fused;
    a.chan_stride = chan_stride;
    a.h = fh;
    a.w = fw;
    a.up = up;
    a.dst = out_hwc;
    const int Wo = fw * up, Ho = fh * up;
    dim3 grid((Wo + kUpTileX - 1) / kUpTileX, (Ho + up - 1) / up, batch);
    const size_t smem = (size_t)kLimbs * kUpMaxRows * (kUpTileX / up + 6) * sizeof(float);
    if (smem > kMaxDynSmem) return fail(IBP_ERR_SHAPE, "ibp_upsample_limbs: up=%d not supported", up);
    StageTimer t(h, IBP_STAGE_UPSAMPLE, (cudaStream_t)stream);
    upsample_limbs_kernel<<<grid, kUpThreads, smem, (cudaStream_t)stream>>>(a);
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

// ------------------------------------------------------------------------------------------------ K2
namespace {
// src: element (img, chan, y, x) at src + img*image_stride + (chan*fh + y)*fw + x; keypoint j in channel chan0 + j;
// twin_base >= 0: average with the mirrored plane twin_base + 30 + flip_heat_ord[j] of the same image on the way in
int launch_peaks(ibp_handle *h, const void *src, int dtype, size_t image_stride, int chan0, int twin_base, int batch, int fh, int fw,
                 int flavour, cudaStream_t st, bool may_defer = false) {
    h->defer_bands = 0;
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "find_peaks: batch %d exceeds limit %d", batch, h->L.max_batch);
    if (fh < 1 || fw < 1 || fw > 32767 || fh > 32767) return fail(IBP_ERR_SHAPE, "find_peaks: bad shape");
    if (flavour != IBP_PEAKS_CROSS_REFINE && flavour != IBP_PEAKS_BOX_CENTROID && flavour != IBP_PEAKS_CROSS_RAW)
        return fail(IBP_ERR_ARG, "find_peaks: flavour %d", flavour);
    const int raw = flavour == IBP_PEAKS_CROSS_RAW;
    if (raw) flavour = IBP_PEAKS_CROSS_REFINE;  // same kernel, the refine step reports the pixel itself
    if (flavour == IBP_PEAKS_BOX_CENTROID && h->P.refine_window != 2)
        return fail(IBP_ERR_ARG, "find_peaks: the centroid flavour supports offset_radius == 2 only");
    PeakArgs a;
    a.src = src;
    a.dtype = dtype;
    a.image_stride = image_stride;
    a.chan0 = chan0;
    a.twin_base = twin_base;
    a.h = fh;
    a.w = fw;
    a.halo = std::max(h->P.refine_window, 1);
    static const bool no_stream = getenv("IBP_PEAKS_NO_STREAM") != nullptr;
    // streaming kernel: rows that the bulk-copy engine can move (16-byte multiples at 16-byte aligned addresses); float16
    // network output (the reference's amp O1 default, parse_skeletons.py:75) streams for the heatmap_nms flavour
    const size_t esz = dtype == IBP_F16 ? 2 : 4;
    const bool stream_ok = ((size_t)fw * esz) % 16 == 0 && ((uintptr_t)src & 15u) == 0 && (image_stride * esz) % 16 == 0 && !no_stream &&
                           (dtype == IBP_F32 || flavour == IBP_PEAKS_CROSS_REFINE);
    // wide maps (image resolution, the original route): bands of at least 16 rows so that the 2*halo rows staged around a band stay a
    // small part of it, as long as two CTAs still fit an SM (40 KB per buffer set; measured on 640x480: 0.425 -> 0.39 ms per 64 images)
    size_t band_target = peak_band_target(stream_ok);
    // the streaming kernel stages the element type of the network output: a float16 band of the same bytes has twice the rows
    // (measured on configs[1] in float16: 64-row bands 651 k img/s against 621 k with 43-row bands)
    const size_t band_esz = stream_ok ? esz : 4;
    if (stream_ok && band_target == kPeakBandTargetStream) band_target = std::max(band_target, std::min<size_t>(40 * 1024, (size_t)16 * fw * band_esz));
    int rows_target = std::max<int>(4 * a.halo, (int)(band_target / ((size_t)fw * band_esz)));
    int nbands = (fh + rows_target - 1) / rows_target;
    a.rows_per_band = (fh + nbands - 1) / nbands;
    a.nbands = (fh + a.rows_per_band - 1) / a.rows_per_band;
    const int rows_buf = std::min(a.rows_per_band + 2 * a.halo, fh);
    size_t band_bytes = std::max((size_t)rows_buf * fw * 4, (size_t)(kParts * a.nbands + 1) * 4);
    band_bytes = std::max(band_bytes, (size_t)batch * 4);
    band_bytes = (band_bytes + 15) & ~(size_t)15;
    a.band_bytes = (int)band_bytes;
    a.up = h->P.upsample;
    a.win = h->P.refine_window;
    a.thr = h->P.peak_thresh;
    a.cap = h->L.max_peaks_per_part;
    a.flavour = flavour;
    a.raw = raw;
    int rc = ensure_bands(h, a.nbands);
    if (rc != IBP_OK) return rc;
    a.band_peaks = h->d_band_peaks;
    a.band_subpix = (flavour == IBP_PEAKS_BOX_CENTROID) ? h->d_band_subpix : nullptr;
    a.band_count = h->d_band_count;
    a.ticket = h->d_ticket;
    a.peaks = h->d_peaks;
    a.subpix = h->d_subpix;
    a.part_count = h->d_part_count;
    a.line = h->d_line;
    a.line_stride = h->line_stride;
    a.n_peaks = h->d_n_peaks;
    a.status = h->d_status;
    a.batch_ticket = h->d_ticket + 2 * h->L.max_batch;
    a.order = h->d_order;
    a.batch = batch;
    a.defer_pack = 0;
    a.yext = nullptr;
    a.n_raw = nullptr;
    a.work_counter = nullptr;
    a.work_base = 0;
    h->order_batch = batch;
    h->use_score64 = 0;
    a.timeline = nullptr;
    {
        const char *dbg = getenv("IBP_DEBUG");
        if (dbg && (atoi(dbg) & 16)) {  // experiments only: per-CTA %globaltimer stamps of the peak kernel
            const size_t n = (size_t)h->L.max_batch * kParts * a.nbands * 6;
            if (!h->d_timeline_pk) {
                CU(cudaMalloc(&h->d_timeline_pk, n * sizeof(unsigned long long)));
                h->timeline_pk_n = n;
            }
            if (h->timeline_pk_n >= n) {
                CU(cudaMemsetAsync(h->d_timeline_pk, 0, n * sizeof(unsigned long long), st));
                a.timeline = h->d_timeline_pk;
            }
        }
    }
    const int side = 2 * a.win + 1;
    size_t scratch = (((size_t)(kPeakThreads / 32) * side * side * a.up * 4) + 7) & ~(size_t)7;
    scratch += (size_t)a.cap * 8;
    scratch = (scratch + 15) & ~(size_t)15;
    CU(cudaMemsetAsync(h->d_status, 0, (size_t)3 * h->L.max_batch * sizeof(uint32_t), st));  // status + raw counts + limb tickets

    // Streaming kernel: float32 planes whose rows the bulk-copy engine can move (16-byte aligned everywhere).
    const bool twin = twin_base >= 0;
    const size_t row_bytes = (size_t)rows_buf * fw * esz;
    const size_t row_buf = (row_bytes + 15) & ~(size_t)15;
    const size_t tail_ints = std::max<size_t>((size_t)kParts * a.nbands + 1, (size_t)batch);
    const size_t smem_stream = 2 * (twin ? 2 : 1) * row_buf + scratch + ((tail_ints * 4 + 15) & ~(size_t)15);
    if (stream_ok && smem_stream <= kMaxDynSmem) {
        a.band_bytes = (int)row_buf;
        a.ring_bytes = 0;
        const int items = a.nbands * kParts * batch;
        void (*kern)(PeakArgs) = nullptr;
        if (dtype == IBP_F16)
            kern = twin ? peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, true, __half> : peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, false, __half>;
        else if (flavour == IBP_PEAKS_CROSS_REFINE)
            kern = twin ? peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, true> : peaks_stream_kernel<IBP_PEAKS_CROSS_REFINE, false>;
        else
            kern = twin ? peaks_stream_kernel<IBP_PEAKS_BOX_CENTROID, true> : peaks_stream_kernel<IBP_PEAKS_BOX_CENTROID, false>;
        if (h->pk_occ_kern != (const void *)kern || h->pk_occ_smem != smem_stream) {
            int per_sm = 0;
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kPeakThreads, smem_stream));
            h->pk_occ_kern = (const void *)kern;
            h->pk_occ_smem = smem_stream;
            h->pk_occ = per_sm;
            const char *dbg = getenv("IBP_DEBUG");
            if (dbg && (atoi(dbg) & 32))
                fprintf(stderr, "[ibp] peaks stream: %d bands of %d rows, %zu B shared, %d CTAs/SM x %d SMs, %d items\n", a.nbands,
                        a.rows_per_band, smem_stream, per_sm, h->sm_count, items);
        }
        const int per_sm = h->pk_occ;
        if (per_sm < 1) return fail(IBP_ERR_SHAPE, "find_peaks: streaming kernel does not fit");
        const int grid = std::min(items, per_sm * h->sm_count);
        // straight into the limb kernel: leave the band lists as they are, the limb CTAs gather them (no tickets, no
        // pack, no fence in this kernel)
        static const bool no_defer = getenv("IBP_PEAKS_NO_DEFER") != nullptr;
        a.defer_pack = (may_defer && a.nbands <= kMaxDeferBands && !no_defer) ? 1 : 0;
        a.n_raw = h->d_n_raw;
        a.yext = a.defer_pack ? h->d_yext : nullptr;
        if (a.defer_pack) {
            h->defer_bands = a.nbands;
            h->defer_batch = batch;
            h->order_batch = 0;
        }
        a.work_counter = h->d_pk_work;
        a.work_base = h->pk_work_base;
        StageTimer t(h, IBP_STAGE_PEAKS, st);
        kern<<<grid, kPeakThreads, smem_stream, st>>>(a);
        CU(cudaGetLastError());
        h->pk_work_base += (unsigned)items;  // once it runs: (items - grid) claims that succeed + one that fails per CTA
        h->launches++;
        if (a.defer_pack) {
            IndexArgs ia;
            ia.band_count = h->d_band_count;
            ia.n_raw = h->d_n_raw;
            ia.nbands = a.nbands;
            ia.batch = batch;
            ia.cap = a.cap;
            ia.line_stride = h->line_stride;
            ia.part_count = h->d_part_count;
            ia.partbase = h->d_partbase;
            ia.n_peaks = h->d_n_peaks;
            ia.order = h->d_order;
            ia.item_meta = h->d_item_meta;
            ia.yext = h->d_yext;
            ia.h = fh;
            ia.w = fw;
            ia.up = h->P.upsample;
            ia.status = h->d_status;
            {
                static const int lag_env = getenv("IBP_GROUP_LAG") ? std::max(0, atoi(getenv("IBP_GROUP_LAG"))) : -1;  // experiments
                ia.group_lag = lag_env >= 0 ? lag_env : kGroupLag;
            }
            CU(launch_dependent(index_bands_kernel, dim3(batch), dim3(32), 0, st, ia));
            h->launches++;
            h->order_batch = batch;
        }
        CU(cudaGetLastError());
        return IBP_OK;
    }

    size_t smem = band_bytes + scratch;
    if (smem > kMaxDynSmem) return fail(IBP_ERR_SHAPE, "find_peaks: rows of %d floats do not fit shared memory", fw);
    a.ring_bytes = 0;
    if (twin_base >= 0 && dtype == IBP_F32 && fw % 4 == 0 && smem + band_bytes <= kMaxDynSmem) {
        a.ring_bytes = (int)band_bytes;  // second band buffer: the twin rows land here by bulk copy
        smem += band_bytes;
    }
    {
        StageTimer t(h, IBP_STAGE_PEAKS, st);
        const dim3 grid(a.nbands, kParts, batch);
        if (flavour == IBP_PEAKS_CROSS_REFINE) {
            if (dtype == IBP_F16)
                peaks_kernel<__half, IBP_PEAKS_CROSS_REFINE><<<grid, kPeakThreads, smem, st>>>(a);
            else
                peaks_kernel<float, IBP_PEAKS_CROSS_REFINE><<<grid, kPeakThreads, smem, st>>>(a);
        } else {
            if (dtype == IBP_F16)
                peaks_kernel<__half, IBP_PEAKS_BOX_CENTROID><<<grid, kPeakThreads, smem, st>>>(a);
            else
                peaks_kernel<float, IBP_PEAKS_BOX_CENTROID><<<grid, kPeakThreads, smem, st>>>(a);
        }
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}
}  // namespace

int ibp_find_peaks(ibp_handle *h, const float *maps, int chan_stride, int first_chan, int batch, int fh, int fw, int flavour, void *stream) {
    if (!h || !maps) return fail(IBP_ERR_ARG, "ibp_find_peaks: null argument");
    if (fh < 1 || fw < 1 || first_chan < 0 || first_chan + kParts > chan_stride) return fail(IBP_ERR_SHAPE, "ibp_find_peaks: bad shape");
    CU(cudaSetDevice(h->device));
    return launch_peaks(h, maps, IBP_F32, (size_t)chan_stride * fh * fw, first_chan, -1, batch, fh, fw, flavour, (cudaStream_t)stream);
}

namespace {
int set_peaks_impl(ibp_handle *h, int image, const float *host_peaks, int n, int row_stride, cudaStream_t st, bool wait);
}
int ibp_set_peaks(ibp_handle *h, int image, const float *host_peaks, int n, int row_stride) {
    return set_peaks_impl(h, image, host_peaks, n, row_stride, nullptr, true);
}
namespace {
// Copies go out asynchronously on `st` from the handle's pinned mirrors; wait = false leaves them in flight (the caller
// launches on the same stream and does not touch the mirrors before it has synchronised).
int set_peaks_impl(ibp_handle *h, int image, const float *host_peaks, int n, int row_stride, cudaStream_t st, bool wait) {
    if (!h || (n > 0 && !host_peaks)) return fail(IBP_ERR_ARG, "ibp_set_peaks: null argument");
    if (image < 0 || image >= h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_set_peaks: image %d out of range", image);
    if (n < 0 || row_stride < 5) return fail(IBP_ERR_SHAPE, "ibp_set_peaks: need rows of >= 5 floats [x,y,score,id,part]");
    CU(cudaSetDevice(h->device));
    const int cap = h->L.max_peaks_per_part;
    int counts[kParts] = {0};
    // pafprocess.cpp:29-41: bucket by part in arrival order; id = arrival index; x,y truncated to int
    for (int i = 0; i < n; i++) {
        const float *r = host_peaks + (size_t)i * row_stride;
        const int part = (int)r[4];
        if (part < 0 || part >= kParts) return fail(IBP_ERR_PART, "joint %d has part id %d outside [0,18)", i, part);
        if (counts[part] >= cap) return fail(IBP_ERR_CAPACITY, "more than %d peaks of part %d (ibp_limits.max_peaks_per_part)", cap, part);
        DevPeak p;
        p.x = (int)r[0];
        p.y = (int)r[1];
        p.score = r[2];
        p.id = i;
        h->h_peaks_stage[(size_t)part * cap + counts[part]++] = p;
    }
    // pafprocess.cpp:43-48: part-major concatenation, indexed later BY PEAK ID (identical to arrival order only when
    // the caller sorted the list by part, as evaluate.py:99-103 does)
    ibp_peak *line = h->h_line + (size_t)image * h->line_stride;
    int pos = 0;
    for (int p = 0; p < kParts; p++)
        for (int k = 0; k < counts[p]; k++) {
            const DevPeak &s = h->h_peaks_stage[(size_t)p * cap + k];
            line[pos].x = s.x;
            line[pos].y = s.y;
            line[pos].score = s.score;
            line[pos].part = p;
            pos++;
        }
    h->h_n_peaks[image] = n;
    h->order_batch = 0;  // the peak kernel's image order no longer describes the lists
    h->defer_bands = 0;
    memcpy(h->h_part_count + (size_t)image * kParts, counts, sizeof(counts));
    // one copy while the per-part stride is small; per part only the filled entries once a handle has grown for a crowd
    if ((size_t)kParts * cap * sizeof(DevPeak) <= 128 * 1024) {
        CU(cudaMemcpyAsync(h->d_peaks + (size_t)image * kParts * cap, h->h_peaks_stage, (size_t)kParts * cap * sizeof(DevPeak),
                           cudaMemcpyHostToDevice, st));
    } else {
        for (int p = 0; p < kParts; p++)
            if (counts[p])
                CU(cudaMemcpyAsync(h->d_peaks + ((size_t)image * kParts + p) * cap, h->h_peaks_stage + (size_t)p * cap,
                                   (size_t)counts[p] * sizeof(DevPeak), cudaMemcpyHostToDevice, st));
    }
    CU(cudaMemcpyAsync(h->d_part_count + (size_t)image * kParts, h->h_part_count + (size_t)image * kParts, sizeof(counts),
                       cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->d_line + (size_t)image * h->line_stride, line, (size_t)std::max(n, 1) * sizeof(ibp_peak), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->d_n_peaks + image, h->h_n_peaks + image, sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(h->d_status + image, 0, sizeof(uint32_t), st));
    if (wait) CU(cudaStreamSynchronize(st));
    return IBP_OK;
}
}  // namespace

// ------------------------------------------------------------------------------------------------ K3 + K4
namespace {
int launch_limbs(ibp_handle *h, const void *limb_src, int dtype, int mode, int twin_base, size_t image_stride, int batch, int dim0, int dim1,
                 int dim2, int min_img_size, cudaStream_t st, bool src_over_bus = false) {
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "assign: batch %d exceeds limit %d", batch, h->L.max_batch);
    if (dim0 < 1 || dim1 < 1 || dim2 < kLimbs) return fail(IBP_ERR_SHAPE, "assign: bad shape (%d,%d,%d)", dim0, dim1, dim2);
    if (mode != IBP_LIMB_FULLRES_HWC && mode != IBP_LIMB_FEAT_PLANAR) return fail(IBP_ERR_ARG, "assign: mode %d", mode);
    LimbArgs a;
    a.peaks = h->d_peaks;
    a.part_count = h->d_part_count;
    a.cap = h->L.max_peaks_per_part;
    a.src = limb_src;
    a.dtype = dtype;
    a.mode = mode;
    a.twin_base = twin_base;
    a.image_stride = image_stride;
    a.up = h->P.upsample;
    size_t plane_bytes = 0;
    if (mode == IBP_LIMB_FULLRES_HWC) {
        a.H = dim0;
        a.W = dim1;
        a.C = dim2;
        a.h = a.w = 1;
        a.stage_plane = 0;
    } else {
        a.h = dim0;
        a.w = dim1;
        a.H = dim0 * a.up;
        a.W = dim1 * a.up;
        a.C = dim2;
        plane_bytes = (size_t)dim0 * dim1 * 4;
        a.stage_plane = (plane_bytes <= kLimbPlaneMaxBytes && limb_smem_bytes(h, plane_bytes) <= kMaxDynSmem) ? 1 : 0;
        if (!a.stage_plane) {
            if (dtype != IBP_F32 || twin_base >= 0)
                return fail(IBP_ERR_SHAPE, "assign: a %dx%d limb plane does not fit shared memory; fuse to float32 first (ibp_fuse_flip)", dim0, dim1);
            plane_bytes = 0;
        }
    }
    a.min_img_size = min_img_size;
    a.thr = h->P.limb_sample_thresh;
    a.step_ratio = h->P.step_ratio;
    a.w_limb = h->P.weight_limb;
    a.w_a = h->P.weight_a;
    a.w_b = h->P.weight_b;
    a.max_steps = h->P.max_steps;
    a.cand_cap = h->L.max_candidates_per_limb;
    a.line = h->d_line;
    a.line_stride = h->line_stride;
    a.links = h->d_links;
    a.link_count = h->d_link_count;
    a.status = h->d_status;
    a.ticket = h->d_ticket + h->L.max_batch;
    a.order = (h->order_batch == batch) ? h->d_order : nullptr;
    a.band_peaks = nullptr;
    a.band_count = nullptr;
    a.nbands = 0;
    a.partbase = nullptr;
    a.item_meta = nullptr;
    a.work_counter = nullptr;
    a.work_base = 0;
    a.plane_bytes = 0;
    a.batch = batch;
    a.peaks_out = h->d_peaks;
    a.line_out = h->d_line;
    a.cta_scratch = nullptr;
    a.cta_scratch_bytes = 0;
    if (h->defer_bands > 0) {
        if (h->defer_batch != batch) return fail(IBP_ERR_ARG, "assign: %d images, but the pending peak lists are of %d", batch, h->defer_batch);
        a.band_peaks = h->d_band_peaks;
        a.band_count = h->d_band_count;
        a.nbands = h->defer_bands;
        a.partbase = h->d_partbase;
        h->defer_bands = 0;
    }
    const size_t smem = limb_smem_bytes(h, plane_bytes);
    a.smem_bytes = (int)smem;
    {
        const char *dbg = getenv("IBP_DEBUG");
        a.debug = dbg ? atoi(dbg) : 0;
        a.timeline = nullptr;
        if (a.debug & 8) {  // experiments only: per-CTA %globaltimer stamps, dumped by ibp_fetch
            if (!h->d_timeline) CU(cudaMalloc(&h->d_timeline, (size_t)h->L.max_batch * 31 * 4 * sizeof(unsigned long long)));
            CU(cudaMemsetAsync(h->d_timeline, 0, (size_t)h->L.max_batch * 31 * 4 * sizeof(unsigned long long), st));
            a.timeline = h->d_timeline;
        }
    }
    GroupArgs &g = a.group;
    g.links = h->d_links;
    g.link_count = h->d_link_count;
    g.cap = a.cap;
    g.max_bodies = h->L.max_bodies;
    g.max_humans = h->L.max_humans;
    g.len_rate = h->P.len_rate;
    g.merge_tol = h->P.merge_tolerance;
    g.min_parts = h->P.min_parts;
    g.min_score = h->P.min_human_score;
    g.humans = h->d_humans;
    g.n_humans = h->d_n_humans;
    g.status = h->d_status;
    // Streaming kernel: the pipeline's common case (see k_limbs.cuh).  The grouping tables must fit behind the plane.
    static const bool no_lstream = getenv("IBP_LIMBS_NO_STREAM") != nullptr;
    const size_t scratch_tail = (size_t)kPairChunk * sizeof(PairGeo) + (size_t)kPairChunk * a.max_steps * 4;
    // the twin rows stream through kLimbRing slots carved from the candidate / geometry / sample scratch: a slot must hold
    // at least one feature row (wider maps take the one-CTA-per-item kernel, whose staging falls back to plain loads)
    const size_t ring_slot_bytes = ((((size_t)a.cand_cap * 16 + scratch_tail) & ~(size_t)15) / kLimbRing);
    // the streaming kernel keeps the plane in the element type of the network output (a float16 flip average is exact in
    // float16): half the shared memory per item for the reference's amp O1 default
    const size_t esz = dtype == IBP_F16 ? 2 : 4;
    const size_t plane_s_bytes = (((size_t)dim0 * dim1 * esz) + 15) & ~(size_t)15;
    const size_t smem_s = limb_smem_bytes(h, plane_s_bytes);
    if (a.band_peaks && mode == IBP_LIMB_FEAT_PLANAR && a.stage_plane && twin_base >= 0 && ((size_t)dim1 * esz) % 16 == 0 &&
        ((uintptr_t)limb_src & 15u) == 0 && (image_stride * esz) % 16 == 0 && h->L.max_peaks_per_part <= 128 &&
        limb_phase_bytes(h, plane_s_bytes) >= group_phase_bytes(h) && smem_s - plane_s_bytes >= group_phase_bytes(h) &&
        scratch_tail >= (size_t)a.cand_cap * 8 && ring_slot_bytes >= (size_t)dim1 * esz && smem_s <= kMaxDynSmem && !no_lstream) {
        a.item_meta = h->d_item_meta;
        a.ticket = h->d_status + 2 * (size_t)h->L.max_batch;  // zeroed by launch_peaks' memset on this stream
        a.work_counter = h->d_lb_work;
        a.work_base = h->lb_work_base;
        a.plane_bytes = (int)plane_s_bytes;
        a.smem_bytes = (int)smem_s;
        const int stream_threads = dtype == IBP_F16 ? kLimbStreamThreadsHalf : kLimbStreamThreads;
        void (*kern)(LimbArgs) = src_over_bus ? (dtype == IBP_F16 ? limbs_stream_kernel<__half, true, kLimbStreamThreadsHalf> : limbs_stream_kernel<float, true>)
                                              : (dtype == IBP_F16 ? limbs_stream_kernel<__half, false, kLimbStreamThreadsHalf> : limbs_stream_kernel<float>);
        const int items = batch * (kLimbs + 1);  // 30 limbs + the grouping of each image
        // Device-resident source: the three-CTAs-per-SM form of k_limbs3.cuh (664 k against 606 k img/s on configs[1],
        // DESIGN.md section 5).  IBP_LIMBS_NO_STREAM3 (read once per process) keeps the two-CTA kernel, which host sources use anyway.
        static const bool no_l3 = getenv("IBP_LIMBS_NO_STREAM3") != nullptr;
        const size_t smem_3 = plane_s_bytes + l3_scratch_bytes(a.cap, a.max_steps);
        if (!src_over_bus && !no_l3 && h->d_l3_scratch && smem_3 >= group_phase_bytes(h) && smem_3 <= kMaxDynSmem &&
            l3_fallback_ring_bytes(a.max_steps) / kLimbRing >= (size_t)dim1 * esz) {
            void (*kern3)(LimbArgs) = dtype == IBP_F16 ? limbs_stream3_kernel<__half> : limbs_stream3_kernel<float>;
            if (h->l3_occ_smem != smem_3 || h->l3_occ_kern != (const void *)kern3) {
                int per_sm = 0;
                CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern3, kL3Threads, smem_3));
                h->l3_occ = std::min(per_sm, kL3MaxCtasPerSm);
                if (const char *cap_env = getenv("IBP_L3_OCC")) h->l3_occ = std::min(h->l3_occ, std::max(atoi(cap_env), 1));  // experiments
                h->l3_occ_smem = smem_3;
                h->l3_occ_kern = (const void *)kern3;
            }
            if (h->l3_occ >= 3 || getenv("IBP_L3_OCC")) {
                a.smem_bytes = (int)smem_3;
                a.cta_scratch = h->d_l3_scratch;
                a.cta_scratch_bytes = l3_global_scratch_bytes(a.cand_cap);
                const int grid = std::min(items, h->l3_occ * h->sm_count);
                StageTimer t(h, IBP_STAGE_LIMBS, st);
                CU(launch_dependent(kern3, dim3(grid), dim3(kL3Threads), smem_3, st, a));
                h->lb_work_base += (unsigned)items;
                h->launches++;
                CU(cudaGetLastError());
                return IBP_OK;
            }
        }
        if (h->lb_occ_smem != smem_s || h->lb_occ_kern != (const void *)kern) {  // every CTA of the grid must be resident (queue items wait for one another)
            int per_sm = 0;
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, stream_threads, smem_s));
            h->lb_occ = per_sm;
            h->lb_occ_smem = smem_s;
            h->lb_occ_kern = (const void *)kern;
        }
        if (h->lb_occ < 1) return fail(IBP_ERR_SHAPE, "assign: streaming limb kernel does not fit");
        const int grid = std::min(items, h->lb_occ * h->sm_count);
        StageTimer t(h, IBP_STAGE_LIMBS, st);
        CU(launch_dependent(kern, dim3(grid), dim3(stream_threads), smem_s, st, a));
        h->lb_work_base += (unsigned)items;  // once it runs, the kernel advances its counter by exactly this much
        h->launches++;
        CU(cudaGetLastError());
        return IBP_OK;
    }
    {
        StageTimer t(h, IBP_STAGE_LIMBS, st);
        const dim3 grid(kLimbs, batch);
        // a handle sized for crowds (more than 128 peaks per joint type) gets the wider CTA: thousands of pairs per limb
        // want more warps per plane in flight; the usual image (tens of pairs) is quicker with fewer, cheaper barriers
        const bool wide = h->L.max_peaks_per_part > 128;
        if (dtype == IBP_F16) {
            if (wide) limbs_kernel<__half, 512><<<grid, 512, smem, st>>>(a);
            else limbs_kernel<__half, 256><<<grid, 256, smem, st>>>(a);
        } else {
            if (wide) limbs_kernel<float, 512><<<grid, 512, smem, st>>>(a);
            else limbs_kernel<float, 256><<<grid, 256, smem, st>>>(a);
        }
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}
}  // namespace

int ibp_assign(ibp_handle *h, const float *limb_src, int mode, int batch, int dim0, int dim1, int dim2, int min_img_size, void *stream) {
    if (!h || !limb_src) return fail(IBP_ERR_ARG, "ibp_assign: null argument");
    CU(cudaSetDevice(h->device));
    const size_t image_stride = (size_t)dim0 * dim1 * (size_t)std::max(dim2, 0);
    return launch_limbs(h, limb_src, IBP_F32, mode, -1, image_stride, batch, dim0, dim1, dim2, min_img_size, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------ pipeline
int ibp_postprocess_batch(ibp_handle *h, const void *net_out, int on_host, int dtype, int batch, int fh, int fw, int min_img_size, void *stream) {
    if (!h || !net_out) return fail(IBP_ERR_ARG, "ibp_postprocess_batch: null argument");
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_postprocess_batch: batch %d exceeds limit %d", batch, h->L.max_batch);
    if (dtype != IBP_F32 && dtype != IBP_F16) return fail(IBP_ERR_ARG, "ibp_postprocess_batch: dtype %d", dtype);
    if (fh < 1 || fw < 1) return fail(IBP_ERR_SHAPE, "ibp_postprocess_batch: bad shape");
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t image_stride = (size_t)2 * kLayers * fh * fw;
    const size_t elems = (size_t)batch * image_stride;
    const void *dev_in = net_out;
    const void *limb_in = nullptr;  // where the limb kernel reads the limb planes from (default: the same place as the peak kernel)
    const size_t plane_bytes = (size_t)fh * fw * 4;
    const bool direct = plane_bytes <= kLimbPlaneMaxBytes && limb_smem_bytes(h, plane_bytes) <= kMaxDynSmem;
    if (on_host) {
        const size_t bytes = elems * (dtype == IBP_F16 ? 2 : 4);
        int rc = ensure_stage(h, bytes, st);
        if (rc != IBP_OK) return rc;
        const size_t half_bytes = bytes / ((size_t)batch * 2);
        // Pinned (page-locked, device-mapped) host buffers: only the 18 keypoint planes of each half are copied — the peak
        // kernel reads all of them — while the limb kernel pulls the limb rows it needs STRAIGHT from the host buffer with its
        // bulk copies: a limb item touches only the rows between the extreme peaks of its two joint types (~43 % of a plane
        // on the benchmark images), and which rows is only known on the device.  36 % fewer bytes cross the bus on a step
        // that is 98 % bus time.  The caller's buffer must stay unchanged until the batch has been fetched.
        static const bool no_zero_copy = getenv("IBP_NO_ZEROCOPY") != nullptr;
        cudaPointerAttributes pa;
        void *mapped = nullptr;
        // only when the streaming limb kernel will run (it alone stages a box per item; the one-CTA-per-item kernel of the crowd
        // handles stages whole planes and is better fed from device memory)
        const size_t esz_in = dtype == IBP_F16 ? 2 : 4;
        const size_t plane_in = (((size_t)fh * fw * esz_in) + 15) & ~(size_t)15;
        const bool will_stream = h->L.max_peaks_per_part <= 128 && ((size_t)fw * esz_in) % 16 == 0 &&
                                 limb_phase_bytes(h, plane_in) >= group_phase_bytes(h) &&
                                 limb_smem_bytes(h, plane_in) - plane_in >= group_phase_bytes(h) && limb_smem_bytes(h, plane_in) <= kMaxDynSmem;
        if (direct && will_stream && !no_zero_copy && cudaPointerGetAttributes(&pa, net_out) == cudaSuccess && pa.type == cudaMemoryTypeHost &&
            pa.devicePointer)
            mapped = pa.devicePointer;
        else
            cudaGetLastError();  // a pageable pointer is not an error here
        if (mapped) {
            const size_t kp_off = half_bytes / kLayers * kLimbs, kp_bytes = half_bytes / kLayers * kParts;
            CU(cudaMemcpy2DAsync((char *)h->d_stage + kp_off, half_bytes, (const char *)net_out + kp_off, half_bytes, kp_bytes, (size_t)batch * 2,
                                 cudaMemcpyHostToDevice, st));
            limb_in = mapped;
        } else {
            // the path reads the 30 limb and 18 keypoint planes of each half (parse_skeletons.py:137 skips the two background
            // channels): only those cross the bus
            const size_t used_bytes = half_bytes / kLayers * (kLimbs + kParts);
            CU(cudaMemcpy2DAsync(h->d_stage, half_bytes, net_out, half_bytes, used_bytes, (size_t)batch * 2, cudaMemcpyHostToDevice, st));
        }
        dev_in = h->d_stage;
    }
    if (direct) {
        // two kernels, the network output is read once and nothing intermediate is written: the flip average
        // happens while each keypoint band / limb plane is staged in shared memory
        int rc = launch_peaks(h, dev_in, dtype, image_stride, kLimbs, kLayers, batch, fh, fw, IBP_PEAKS_CROSS_REFINE, st, /*may_defer=*/true);
        if (rc != IBP_OK) return rc;
        return launch_limbs(h, limb_in ? limb_in : dev_in, dtype, IBP_LIMB_FEAT_PLANAR, kLayers, image_stride, batch, fh, fw, kLayers, min_img_size, st,
                            /*src_over_bus=*/limb_in != nullptr);
    }
    // limb planes too large for shared memory: materialise the fused float32 maps, sample them through L2
    int rc = ensure_fused(h, (size_t)batch * kLayers * fh * fw);
    if (rc != IBP_OK) return rc;
    rc = ibp_fuse_flip(h, dev_in, dtype, batch, fh, fw, 1, h->d_fused, stream);
    if (rc != IBP_OK) return rc;
    rc = ibp_find_peaks(h, h->d_fused, kLayers, kLimbs, batch, fh, fw, IBP_PEAKS_CROSS_REFINE, stream);
    if (rc != IBP_OK) return rc;
    return ibp_assign(h, h->d_fused, IBP_LIMB_FEAT_PLANAR, batch, fh, fw, kLayers, min_img_size, stream);
}

int ibp_fetch(ibp_handle *h, int batch, ibp_batch_view *view, void *stream) {
    if (!h || !view) return fail(IBP_ERR_ARG, "ibp_fetch: null argument");
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_fetch: batch %d exceeds limit %d", batch, h->L.max_batch);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t B = batch;
    CU(cudaMemcpyAsync(h->h_n_peaks, h->d_n_peaks, B * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_part_count, h->d_part_count, B * kParts * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_line, h->d_line, B * h->line_stride * sizeof(ibp_peak), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_n_humans, h->d_n_humans, B * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_humans, h->d_humans, B * h->L.max_humans * sizeof(ibp_human), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_status, h->d_status, B * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (h->d_timeline) {
        std::vector<unsigned long long> tl((size_t)batch * 31 * 4);
        CU(cudaMemcpy(tl.data(), h->d_timeline, tl.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        const char *path = getenv("IBP_TIMELINE");
        if (FILE *f = fopen(path ? path : "ibp_timeline.bin", "wb")) {
            fwrite(tl.data(), sizeof(unsigned long long), tl.size(), f);
            fclose(f);
        }
    }
    if (h->d_timeline_pk) {
        std::vector<unsigned long long> tl(h->timeline_pk_n);
        CU(cudaMemcpy(tl.data(), h->d_timeline_pk, tl.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        const char *path = getenv("IBP_TIMELINE_PK");
        if (FILE *f = fopen(path ? path : "ibp_timeline_pk.bin", "wb")) {
            fwrite(tl.data(), sizeof(unsigned long long), tl.size(), f);
            fclose(f);
        }
    }
    view->batch = batch;
    view->peak_stride = h->line_stride;
    view->human_stride = h->L.max_humans;
    view->n_peaks = h->h_n_peaks;
    view->part_count = h->h_part_count;
    view->peaks = h->h_line;
    view->n_humans = h->h_n_humans;
    view->humans = h->h_humans;
    view->status = h->h_status;
    uint32_t any = 0;
    for (int b = 0; b < batch; b++) any |= h->h_status[b];
    if (any & IBP_ST_STALLED) return fail(IBP_ERR_CUDA, "internal: the grouping of an image never saw all of its limbs finish");
    if (any & (IBP_ST_PEAK_OVERFLOW | IBP_ST_CAND_OVERFLOW | IBP_ST_BODY_OVERFLOW | IBP_ST_HUMAN_OVERFLOW))
        return fail(IBP_ERR_CAPACITY, "a device buffer overflowed (status bits 0x%x); raise ibp_limits", any);
    return IBP_OK;
}

int ibp_fetch_coco(ibp_handle *h, int batch, ibp_coco_view *view, void *stream) {
    if (!h || !view) return fail(IBP_ERR_ARG, "ibp_fetch_coco: null argument");
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_fetch_coco: batch %d exceeds limit %d", batch, h->L.max_batch);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t B = h->L.max_batch, cap_rows = B * h->L.max_humans;
    if (!h->d_coco_off) {
        CU(cudaMalloc(&h->d_coco_off, (B + 1) * sizeof(int)));
        CU(cudaMalloc(&h->d_coco_rows, cap_rows * kCocoRow * sizeof(float)));
        CU(cudaMallocHost(&h->h_coco_off, (B + 1) * sizeof(int)));
        CU(cudaMallocHost(&h->h_coco_rows, cap_rows * kCocoRow * sizeof(float)));
    }
    CocoArgs a;
    a.n_humans = h->d_n_humans;
    a.humans = h->d_humans;
    a.line = h->d_line;
    a.batch = batch;
    a.human_stride = h->L.max_humans;
    a.line_stride = h->line_stride;
    a.offsets = h->d_coco_off;
    a.rows = h->d_coco_rows;
    coco_rows_kernel<<<batch, 128, 0, st>>>(a);
    h->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(h->h_coco_off, h->d_coco_off, ((size_t)batch + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_status, h->d_status, (size_t)batch * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const size_t total = (size_t)h->h_coco_off[batch];
    if (total) {
        CU(cudaMemcpyAsync(h->h_coco_rows, h->d_coco_rows, total * kCocoRow * sizeof(float), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    view->batch = batch;
    view->offsets = h->h_coco_off;
    view->rows = h->h_coco_rows;
    view->status = h->h_status;
    uint32_t any = 0;
    for (int b = 0; b < batch; b++) any |= h->h_status[b];
    if (any & IBP_ST_STALLED) return fail(IBP_ERR_CUDA, "internal: the grouping of an image never saw all of its limbs finish");
    if (any & (IBP_ST_PEAK_OVERFLOW | IBP_ST_CAND_OVERFLOW | IBP_ST_BODY_OVERFLOW | IBP_ST_HUMAN_OVERFLOW))
        return fail(IBP_ERR_CAPACITY, "a device buffer overflowed (status bits 0x%x); raise ibp_limits", any);
    return IBP_OK;
}

int ibp_prepare_image(ibp_handle *h, const uint8_t *img, int batch, int img_h, int img_w, int max_downsample, int pad_value, int flip,
                      float *out, void *stream) {
    if (!h || !img || !out) return fail(IBP_ERR_ARG, "ibp_prepare_image: null argument");
    if (batch < 1 || img_h < 1 || img_w < 1 || max_downsample < 1) return fail(IBP_ERR_SHAPE, "ibp_prepare_image: bad shape");
    if (pad_value < 0 || pad_value > 255) return fail(IBP_ERR_ARG, "ibp_prepare_image: pad value %d is not a uint8", pad_value);
    if (batch > 65535) return fail(IBP_ERR_ARG, "ibp_prepare_image: batch %d too large for one launch", batch);
    CU(cudaSetDevice(h->device));
    PrepArgs a;
    a.img = img;
    a.out = out;
    a.batch = batch;
    a.h = img_h;
    a.w = img_w;
    // utils/util.py:59-62 tiles row -2 / column -2 of the image: empty for a one-row / one-column image, which stays unpadded
    a.hp = img_h == 1 ? 1 : (img_h + max_downsample - 1) / max_downsample * max_downsample;
    a.wp = img_w == 1 ? 1 : (img_w + max_downsample - 1) / max_downsample * max_downsample;
    if (a.hp > 65535) return fail(IBP_ERR_SHAPE, "ibp_prepare_image: image too tall (%d rows)", a.hp);
    a.flip = flip ? 1 : 0;
    a.pad_value = pad_value;
    const dim3 grid((a.wp + 255) / 256, a.hp, batch);
    prepare_image_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

void ibp_resized_size(int img_h, int img_w, double fx, double fy, int *out_h, int *out_w) {
    // cv2.resize with dsize = (0, 0): Size(saturate_cast<int>(cols * fx), saturate_cast<int>(rows * fy)), which rounds half to even
    if (out_h) *out_h = (int)nearbyint((double)img_h * fy);
    if (out_w) *out_w = (int)nearbyint((double)img_w * fx);
}

int ibp_resize_image(ibp_handle *h, const uint8_t *img, int batch, int img_h, int img_w, double fx, double fy, uint8_t *out, int out_h,
                     int out_w, void *stream) {
    if (!h || !img || !out) return fail(IBP_ERR_ARG, "ibp_resize_image: null argument");
    if (batch < 1 || batch > 65535 || img_h < 1 || img_w < 1 || !(fx > 0.0) || !(fy > 0.0)) return fail(IBP_ERR_SHAPE, "ibp_resize_image: bad shape / factor");
    int oh = 0, ow = 0;
    ibp_resized_size(img_h, img_w, fx, fy, &oh, &ow);
    if (oh < 1 || ow < 1 || oh > 65535) return fail(IBP_ERR_SHAPE, "ibp_resize_image: a %dx%d image times (%g, %g) gives %dx%d", img_h, img_w, fy, fx, oh, ow);
    if (oh != out_h || ow != out_w)
        return fail(IBP_ERR_SHAPE, "ibp_resize_image: cv2 would produce %dx%d for these factors, the output buffer is %dx%d", oh, ow, out_h, out_w);
    CU(cudaSetDevice(h->device));
    ResizeArgs a;
    a.img = img;
    a.out = out;
    a.batch = batch;
    a.h = img_h;
    a.w = img_w;
    a.dh = oh;
    a.dw = ow;
    a.scale_x = 1.0 / fx;  // dsize empty: scale = 1 / inv_scale, not src / dst
    a.scale_y = 1.0 / fy;
    a.lanes = 8;
    resize_u8_cubic_kernel<<<dim3((ow + 127) / 128, oh, batch), 128, 0, (cudaStream_t)stream>>>(a);
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

void ibp_rotation_matrix(double cx, double cy, double angle_deg, double scale, double *m) {
    // cv::getRotationMatrix2D (imgwarp.cpp): angle *= CV_PI / 180; alpha = cos * scale; beta = sin * scale
    if (!m) return;
    const double angle = angle_deg * (3.1415926535897932384626433832795 / 180);
    const double alpha = std::cos(angle) * scale, beta = std::sin(angle) * scale;
    cx = (double)(float)cx;  // the centre is a Point2f
    cy = (double)(float)cy;
    m[0] = alpha;
    m[1] = beta;
    m[2] = (1 - alpha) * cx - beta * cy;
    m[3] = -beta;
    m[4] = alpha;
    m[5] = beta * cx + (1 - alpha) * cy;
}

int ibp_warp_affine(ibp_handle *h, const float *src, float *dst, int batch, int channels, int hh, int ww, int64_t batch_stride,
                    int64_t chan_stride, int64_t row_stride, int64_t pix_stride, const double *m6, void *stream) {
    if (!h || !src || !dst || !m6) return fail(IBP_ERR_ARG, "ibp_warp_affine: null argument");
    if (src == dst) return fail(IBP_ERR_ARG, "ibp_warp_affine: in-place warp is not supported");
    if (batch < 1 || batch > 65535 || channels < 1 || hh < 1 || ww < 1 || hh > 65535) return fail(IBP_ERR_SHAPE, "ibp_warp_affine: bad shape");
    CU(cudaSetDevice(h->device));
    WarpArgs a;
    a.src = src;
    a.dst = dst;
    a.batch = batch;
    a.channels = channels;
    a.h = hh;
    a.w = ww;
    a.bs = batch_stride;
    a.cs = chan_stride;
    a.ys = row_stride;
    a.xs = pix_stride;
    // cv::warpAffine without WARP_INVERSE_MAP: invert the 2x3 matrix in double, operation by operation
    double M[6];
    for (int i = 0; i < 6; i++) M[i] = m6[i];
    double D = M[0] * M[4] - M[1] * M[3];
    D = D != 0 ? 1. / D : 0;
    const double A11 = M[4] * D, A22 = M[0] * D;
    M[0] = A11;
    M[1] *= -D;
    M[3] *= -D;
    M[4] = A22;
    const double b1 = -M[0] * M[2] - M[1] * M[5];
    const double b2 = -M[3] * M[2] - M[4] * M[5];
    M[2] = b1;
    M[5] = b2;
    for (int i = 0; i < 6; i++) a.m[i] = M[i];
    warp_affine_kernel<<<dim3((ww + 127) / 128, hh, batch), 128, 0, (cudaStream_t)stream>>>(a);
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

int ibp_crop_resize_accumulate(ibp_handle *h, const float *maps, int batch, int ph, int pw, int sh, int sw, int img_h, int img_w, int n_items,
                               double *out, void *stream) {
    if (!h || !maps || !out) return fail(IBP_ERR_ARG, "ibp_crop_resize_accumulate: null argument");
    if (batch < 1 || batch > 65535 || ph < 1 || pw < 1 || sh < 1 || sw < 1 || sh > ph || sw > pw || img_h < 1 || img_w < 1 || img_h > 65535 || n_items < 1)
        return fail(IBP_ERR_SHAPE, "ibp_crop_resize_accumulate: bad shape (maps %dx%d, crop %dx%d, image %dx%d, %d items)", ph, pw, sh, sw, img_h, img_w, n_items);
    CU(cudaSetDevice(h->device));
    CropResizeArgs a;
    a.src = maps;
    a.batch = batch;
    a.ph = ph;
    a.pw = pw;
    a.sh = sh;
    a.sw = sw;
    a.H = img_h;
    a.W = img_w;
    a.scale_x = 1.0 / ((double)img_w / (double)sw);
    a.scale_y = 1.0 / ((double)img_h / (double)sh);
    a.divisor = (float)n_items;
    a.out = out;
    crop_resize_accumulate_kernel<<<dim3((img_w + 127) / 128, img_h, batch), 128, 0, (cudaStream_t)stream>>>(a);
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

int ibp_fetch_subpix(ibp_handle *h, int image, double *host_xy, int capacity, void *stream) {
    if (!h || !host_xy) return fail(IBP_ERR_ARG, "ibp_fetch_subpix: null argument");
    if (image < 0 || image >= h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_fetch_subpix: image %d out of range", image);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cap = h->L.max_peaks_per_part;
    std::vector<double> tmp(kParts * cap * 2);
    CU(cudaMemcpyAsync(tmp.data(), h->d_subpix + (size_t)image * kParts * cap * 2, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    // per-part lists -> id order (running counter across parts), using the counts of the last ibp_fetch
    int pos = 0;
    for (int p = 0; p < kParts; p++) {
        const int n = h->h_part_count[(size_t)image * kParts + p];
        for (int k = 0; k < n; k++, pos++) {
            if (pos >= capacity) return fail(IBP_ERR_CAPACITY, "ibp_fetch_subpix: more than %d peaks", capacity);
            host_xy[2 * pos] = tmp[((size_t)p * cap + k) * 2];
            host_xy[2 * pos + 1] = tmp[((size_t)p * cap + k) * 2 + 1];
        }
    }
    return pos;
}

// ------------------------------------------------------------------------------------------------ Python route
namespace {
PyParams to_py(const ibp_py_params *p) {
    ibp_py_params d;
    if (!p) {
        ibp_default_py_params(&d);
        p = &d;
    }
    PyParams q;
    q.thre2 = p->thre2;
    q.connect_ration = p->connect_ration;
    q.mid_num = p->mid_num;
    q.len_rate = p->len_rate;
    q.connection_tole = p->connection_tole;
    q.min_score = p->min_score;
    q.min_parts = p->min_parts;
    return q;
}
}  // namespace

void ibp_default_py_params(ibp_py_params *p) {
    if (!p) return;
    p->thre2 = 0.1;
    p->connect_ration = 0.8;
    p->mid_num = 20;
    p->len_rate = 16;
    p->connection_tole = 0.7;
    p->min_score = 0.45;
    p->min_parts = 2;
}

int ibp_set_peaks64(ibp_handle *h, int image, const double *rows, const int32_t *part, int n) {
    if (!h || (n > 0 && (!rows || !part))) return fail(IBP_ERR_ARG, "ibp_set_peaks64: null argument");
    if (image < 0 || image >= h->L.max_batch || n < 0) return fail(IBP_ERR_ARG, "ibp_set_peaks64: image %d / n %d out of range", image, n);
    CU(cudaSetDevice(h->device));
    int rc = ensure_py(h);
    if (rc != IBP_OK) return rc;
    const int cap = h->L.max_peaks_per_part;
    std::vector<DevPeak> pk((size_t)kParts * cap);
    std::vector<double> xy((size_t)kParts * cap * 2), sc((size_t)kParts * cap);
    int counts[kParts] = {0};
    for (int i = 0; i < n; i++) {
        const int p = part[i];
        if (p < 0 || p >= kParts) return fail(IBP_ERR_PART, "joint %d has part id %d outside [0,18)", i, p);
        if (counts[p] >= cap) return fail(IBP_ERR_CAPACITY, "more than %d peaks of part %d (ibp_limits.max_peaks_per_part)", cap, p);
        const size_t k = (size_t)p * cap + counts[p]++;
        const double *r = rows + (size_t)i * 4;
        pk[k].x = (int)r[0];
        pk[k].y = (int)r[1];
        pk[k].score = (float)r[2];
        pk[k].id = (int)r[3];
        xy[2 * k] = r[0];
        xy[2 * k + 1] = r[1];
        sc[k] = r[2];
    }
    const size_t off = (size_t)image * kParts * cap;
    CU(cudaMemcpy(h->d_peaks + off, pk.data(), pk.size() * sizeof(DevPeak), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_subpix + off * 2, xy.data(), xy.size() * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_score64 + off, sc.data(), sc.size() * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_part_count + (size_t)image * kParts, counts, sizeof(counts), cudaMemcpyHostToDevice));
    CU(cudaMemset(h->d_status + image, 0, sizeof(uint32_t)));
    memcpy(h->h_part_count + (size_t)image * kParts, counts, sizeof(counts));
    h->h_n_peaks[image] = n;
    h->use_score64 = 1;
    h->order_batch = 0;
    h->defer_bands = 0;
    return IBP_OK;
}

int ibp_find_connections(ibp_handle *h, const ibp_py_params *params, const double *paf, int chan_stride, int batch, int img_h, int img_w,
                         int img_height, void *stream) {
    if (!h || !paf) return fail(IBP_ERR_ARG, "ibp_find_connections: null argument");
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_find_connections: batch %d exceeds limit %d", batch, h->L.max_batch);
    if (img_h < 1 || img_w < 1 || chan_stride < kLimbs) return fail(IBP_ERR_SHAPE, "ibp_find_connections: bad shape");
    CU(cudaSetDevice(h->device));
    int rc = ensure_py(h);
    if (rc != IBP_OK) return rc;
    PyConnArgs a;
    a.P = to_py(params);
    if (a.P.mid_num < 1 || a.P.mid_num > kPyMaxSteps) return fail(IBP_ERR_ARG, "ibp_find_connections: mid_num must be in [1,%d]", kPyMaxSteps);
    a.peaks = h->d_peaks;
    a.subpix = h->d_subpix;
    a.score64 = h->use_score64 ? h->d_score64 : nullptr;
    a.part_count = h->d_part_count;
    a.cap = h->L.max_peaks_per_part;
    a.paf = paf;
    a.paf_chan_stride = chan_stride;
    a.H = img_h;
    a.W = img_w;
    a.img_height = img_height;
    a.cand_cap = h->L.max_candidates_per_limb;
    a.conn = h->d_py_conn;
    a.conn_count = h->d_py_conn_count;
    a.cand_score = h->d_py_cand;
    a.line_stride = h->line_stride;
    a.status = h->d_status;
    const size_t smem = (size_t)a.cap * (2 * sizeof(PyPeak) + 2) + (size_t)a.cand_cap * (3 * 8 + 4 + 4) + 16;
    if (smem > kMaxDynSmem) return fail(IBP_ERR_ARG, "ibp_find_connections: limits need %zu B of shared memory", smem);
    {
        StageTimer t(h, IBP_STAGE_LIMBS, (cudaStream_t)stream);
        py_connections_kernel<<<dim3(kLimbs, batch), kPyThreads, smem, (cudaStream_t)stream>>>(a);
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

int ibp_fetch_connections(ibp_handle *h, int image, double *rows, int32_t *counts, void *stream) {
    if (!h || !rows || !counts) return fail(IBP_ERR_ARG, "ibp_fetch_connections: null argument");
    if (image < 0 || image >= h->L.max_batch || !h->d_py_conn) return fail(IBP_ERR_ARG, "ibp_fetch_connections: nothing to fetch for image %d", image);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cap = h->L.max_peaks_per_part;
    uint32_t status = 0;
    CU(cudaMemcpyAsync(counts, h->d_py_conn_count + (size_t)image * kLimbs, kLimbs * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(rows, h->d_py_conn + (size_t)image * kLimbs * cap * 6, kLimbs * cap * 6 * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&status, h->d_status + image, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (status & (IBP_ST_PEAK_OVERFLOW | IBP_ST_CAND_OVERFLOW))
        return fail(IBP_ERR_CAPACITY, "a device buffer overflowed (status bits 0x%x); raise ibp_limits", status);
    return IBP_OK;
}

int ibp_set_connections(ibp_handle *h, int image, const double *rows, const int32_t *counts, const double *cand_score, int n_cand) {
    if (!h || !rows || !counts || (n_cand > 0 && !cand_score)) return fail(IBP_ERR_ARG, "ibp_set_connections: null argument");
    if (image < 0 || image >= h->L.max_batch || n_cand < 0) return fail(IBP_ERR_ARG, "ibp_set_connections: image %d out of range", image);
    if (n_cand > h->line_stride) return fail(IBP_ERR_CAPACITY, "ibp_set_connections: %d joint candidates exceed %d", n_cand, h->line_stride);
    CU(cudaSetDevice(h->device));
    int rc = ensure_py(h);
    if (rc != IBP_OK) return rc;
    const size_t cap = h->L.max_peaks_per_part;
    for (int l = 0; l < kLimbs; l++)
        if (counts[l] < 0 || (size_t)counts[l] > cap) return fail(IBP_ERR_CAPACITY, "ibp_set_connections: limb %d has %d rows (max %zu)", l, counts[l], cap);
    CU(cudaMemcpy(h->d_py_conn + (size_t)image * kLimbs * cap * 6, rows, kLimbs * cap * 6 * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_py_conn_count + (size_t)image * kLimbs, counts, kLimbs * sizeof(int), cudaMemcpyHostToDevice));
    if (n_cand > 0) CU(cudaMemcpy(h->d_py_cand + (size_t)image * h->line_stride, cand_score, (size_t)n_cand * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemset(h->d_status + image, 0, sizeof(uint32_t)));
    return IBP_OK;
}

int ibp_find_humans(ibp_handle *h, const ibp_py_params *params, int batch, void *stream) {
    if (!h) return fail(IBP_ERR_ARG, "ibp_find_humans: null handle");
    if (batch < 1 || batch > h->L.max_batch) return fail(IBP_ERR_ARG, "ibp_find_humans: batch %d exceeds limit %d", batch, h->L.max_batch);
    CU(cudaSetDevice(h->device));
    int rc = ensure_py(h);
    if (rc != IBP_OK) return rc;
    PyHumanArgs a;
    a.conn = h->d_py_conn;
    a.conn_count = h->d_py_conn_count;
    a.part_count = h->d_part_count;
    a.cap = h->L.max_peaks_per_part;
    a.cand_score = h->d_py_cand;
    a.line_stride = h->line_stride;
    a.P = to_py(params);
    a.max_bodies = h->L.max_bodies;
    a.max_humans = h->L.max_humans;
    a.people = h->d_py_people;
    a.n_people = h->d_py_n_people;
    a.status = h->d_status;
    const size_t smem = (size_t)a.max_bodies * 40 * sizeof(double);
    if (smem > kMaxDynSmem) return fail(IBP_ERR_ARG, "ibp_find_humans: max_bodies=%d needs %zu B of shared memory (max %zu)", a.max_bodies, smem, kMaxDynSmem);
    {
        StageTimer t(h, IBP_STAGE_GROUP, (cudaStream_t)stream);
        py_humans_kernel<<<batch, 32, smem, (cudaStream_t)stream>>>(a);
    }
    h->launches++;
    CU(cudaGetLastError());
    return IBP_OK;
}

int ibp_fetch_people(ibp_handle *h, int image, double *people, int capacity, void *stream) {
    if (!h || !people) return fail(IBP_ERR_ARG, "ibp_fetch_people: null argument");
    if (image < 0 || image >= h->L.max_batch || !h->d_py_people) return fail(IBP_ERR_ARG, "ibp_fetch_people: nothing to fetch for image %d", image);
    CU(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    uint32_t status = 0;
    CU(cudaMemcpyAsync(&n, h->d_py_n_people + image, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&status, h->d_status + image, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (status & (IBP_ST_PEAK_OVERFLOW | IBP_ST_CAND_OVERFLOW | IBP_ST_BODY_OVERFLOW | IBP_ST_HUMAN_OVERFLOW))
        return fail(IBP_ERR_CAPACITY, "a device buffer overflowed (status bits 0x%x); raise ibp_limits", status);
    if (n > capacity) return fail(IBP_ERR_CAPACITY, "ibp_fetch_people: %d people, room for %d", n, capacity);
    if (n > 0) {
        CU(cudaMemcpyAsync(people, h->d_py_people + (size_t)image * h->L.max_humans * 40, (size_t)n * 40 * sizeof(double), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    return n;
}

// ------------------------------------------------------------------------------------------------ legacy protocol
namespace {
ibp_handle *g_legacy = nullptr;
ibp_batch_view g_view;
bool g_view_valid = false;
uint32_t g_legacy_status = 0;

// Pageable host memory -> device through pinned slots that several host threads fill in parallel: one thread's memcpy
// into a pinned buffer (what a pageable cudaMemcpy does inside the driver) runs at 10-12 GB/s, a fraction of the bus.
// Each worker owns kSlots pinned slots and a stream of its own; chunk c goes to worker c % T.  The consumer stream waits
// for every worker's last DMA through an event.  Workers are detached and live for the process (the legacy protocol is
// process-global state anyway, pafprocess.cpp:16-17).
class HostStager {
public:
    static constexpr size_t kChunk = 512 * 1024;
    static constexpr int kSlots = 3;
    static HostStager *get(int device) {
        static std::mutex mu;
        static HostStager *inst = nullptr;
        std::lock_guard<std::mutex> lk(mu);
        if (!inst || inst->device_ != device) {
            HostStager *s = new HostStager();
            if (!s->init(device)) return nullptr;  // (leaks a few words on failure; the caller falls back to cudaMemcpy)
            inst = s;
        }
        return inst;
    }
    // enqueue the upload of [src, src + bytes) to dst and make `consumer` wait for it
    bool upload(const char *src, char *dst, size_t bytes, cudaStream_t consumer) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            src_ = src;
            dst_ = dst;
            bytes_ = bytes;
            pending_ = nthreads_;
            failed_ = false;
            gen_++;
        }
        cv_.notify_all();
        {
            std::unique_lock<std::mutex> lk(mu_);
            done_cv_.wait(lk, [&] { return pending_ == 0; });
        }
        if (failed_) return false;
        for (int t = 0; t < nthreads_; t++)
            if (cudaStreamWaitEvent(consumer, w_[t].done, 0) != cudaSuccess) return false;
        return true;
    }

private:
    struct Worker {
        char *slot[kSlots] = {nullptr};
        cudaEvent_t free_ev[kSlots] = {nullptr};
        cudaStream_t st = nullptr;
        cudaEvent_t done = nullptr;
    };
    bool init(int device) {
        device_ = device;
        unsigned hc = std::thread::hardware_concurrency();
        nthreads_ = (int)std::max(1u, std::min(4u, hc ? hc : 1u));
        const char *e = getenv("IBP_STAGER_THREADS");
        if (e && atoi(e) >= 1) nthreads_ = std::min(atoi(e), 16);
        w_.resize(nthreads_);
        if (cudaSetDevice(device) != cudaSuccess) return false;
        for (auto &w : w_) {
            if (cudaStreamCreateWithFlags(&w.st, cudaStreamNonBlocking) != cudaSuccess) return false;
            if (cudaEventCreateWithFlags(&w.done, cudaEventDisableTiming) != cudaSuccess) return false;
            for (int k = 0; k < kSlots; k++) {
                if (cudaMallocHost(&w.slot[k], kChunk) != cudaSuccess) return false;
                if (cudaEventCreateWithFlags(&w.free_ev[k], cudaEventDisableTiming) != cudaSuccess) return false;
            }
        }
        for (int t = 0; t < nthreads_; t++) std::thread([this, t] { run(t); }).detach();
        return true;
    }
    void run(int t) {
        cudaSetDevice(device_);
        Worker &w = w_[t];
        unsigned long long seen = 0;
        for (;;) {
            const char *src;
            char *dst;
            size_t bytes;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                src = src_, dst = dst_, bytes = bytes_;
            }
            bool ok = true;
            const size_t nchunks = (bytes + kChunk - 1) / kChunk;
            int k = 0;
            for (size_t c = t; c < nchunks && ok; c += nthreads_, k = (k + 1) % kSlots) {
                const size_t off = c * kChunk, len = std::min(kChunk, bytes - off);
                ok = cudaEventSynchronize(w.free_ev[k]) == cudaSuccess;  // the slot's previous DMA has left it
                if (!ok) break;
                memcpy(w.slot[k], src + off, len);
                ok = cudaMemcpyAsync(dst + off, w.slot[k], len, cudaMemcpyHostToDevice, w.st) == cudaSuccess &&
                     cudaEventRecord(w.free_ev[k], w.st) == cudaSuccess;
            }
            ok = ok && cudaEventRecord(w.done, w.st) == cudaSuccess;
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (!ok) failed_ = true;
                pending_--;
            }
            done_cv_.notify_all();
        }
    }
    int device_ = -1, nthreads_ = 1;
    std::vector<Worker> w_;
    std::mutex mu_;
    std::condition_variable cv_, done_cv_;
    unsigned long long gen_ = 0;
    int pending_ = 0;
    bool failed_ = false;
    const char *src_ = nullptr;
    char *dst_ = nullptr;
    size_t bytes_ = 0;
};
}  // namespace

namespace {
// Legacy entry point, limb map in pinned host memory: sample it in place over the bus instead of copying the rows between
// the extreme peaks?  A sample costs a 32-byte read at a small fraction of the bus's streaming rate (taken as 1/8 here), the
// copy moves the rows at full rate.
bool legacy_sample_over_bus(const cudaPointerAttributes &attr, const int *counts, int max_steps, int f1, int f2, int f3, int y_lo, int y_hi) {
    static const bool off = getenv("IBP_NO_ZEROCOPY") != nullptr;
    if (off || attr.type != cudaMemoryTypeHost || !attr.devicePointer) return false;
    static const int limb_a[kLimbs] = {1, 1, 1, 1, 1, 0, 0, 14, 15, 1, 2, 3, 1, 5, 6, 1, 8, 9, 1, 11, 12, 0, 0, 2, 8, 5, 11, 16, 17, 8};
    static const int limb_b[kLimbs] = {0, 14, 15, 16, 17, 14, 15, 16, 17, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 2, 5, 8, 12, 11, 9, 2, 5, 11};
    size_t samples = 0;
    for (int l = 0; l < kLimbs; l++) samples += (size_t)counts[limb_a[l]] * counts[limb_b[l]] * (size_t)max_steps;
    const int r0 = std::max(0, std::min(y_lo, f1 - 1) - 1), r1 = std::min(f1 - 1, std::max(y_hi, 0) + 1);
    const size_t copy_bytes = (size_t)(r1 - r0 + 1) * f2 * f3 * sizeof(float);
    return samples * 32 * 8 < copy_bytes;
}
}  // namespace

int ibp_process_paf(int p1, int p2, int p3, const float *peaks, int f1, int f2, int f3, const float *pafmap, int min_img_size) {
    g_view_valid = false;
    if (!peaks || !pafmap) return fail(IBP_ERR_ARG, "process_paf: null array");
    if (p1 < 0 || p2 < 0 || p3 < 5 || f1 < 1 || f2 < 1 || f3 < kLimbs)
        return fail(IBP_ERR_SHAPE, "process_paf: peaks must be (p1,p2,>=5) and pafmap (H,W,>=30); got (%d,%d,%d) (%d,%d,%d)", p1, p2, p3, f1, f2, f3);
    const int n = p1 * p2;
    // capacity: grow the process-global handle when a part has more peaks than it can hold; y extent of the joint list
    int counts[kParts] = {0};
    int y_lo = INT_MAX, y_hi = INT_MIN;
    for (int i = 0; i < n; i++) {
        const float *r = peaks + (size_t)i * p3;
        const int part = (int)r[4];
        if (part < 0 || part >= kParts) return fail(IBP_ERR_PART, "joint %d has part id %d outside [0,18)", i, part);
        counts[part]++;
        const int y = (int)r[1];  // pafprocess.cpp:36
        y_lo = std::min(y_lo, y);
        y_hi = std::max(y_hi, y);
    }
    int need = 128;
    for (int p = 0; p < kParts; p++)
        while (need < counts[p]) need <<= 1;
    if (!g_legacy || g_legacy->L.max_peaks_per_part < need) {
        if (g_legacy) ibp_destroy(g_legacy);
        g_legacy = nullptr;
        ibp_limits l;
        ibp_default_limits(&l);
        l.max_batch = 1;
        l.max_peaks_per_part = need;
        l.max_candidates_per_limb = 4096;
        l.max_bodies = 1024;
        l.max_humans = 1024;
        const char *dev = getenv("IBP_DEVICE");
        int rc = ibp_create(&g_legacy, dev ? atoi(dev) : 0, nullptr, &l);
        if (rc != IBP_OK) return rc;
    }
    ibp_handle *h = g_legacy;
    CU(cudaSetDevice(h->device));
    cudaStream_t st = nullptr;  // the legacy default stream: ordered after whatever the caller did on it (a CUDA tensor argument)
    int rc = set_peaks_impl(h, 0, peaks, n, p3, st, false);
    if (rc != IBP_OK) return rc;

    // The limb map.  Where does the caller's array live?  Device memory (a CUDA tensor handed through
    // __cuda_array_interface__): sampled in place.  Host memory: only the image rows between the extreme peaks can be
    // sampled (every sample lies on the segment between two peaks, pafprocess.cpp:319-320; one spare row either side),
    // so only those rows cross the bus — pinned memory straight by DMA, pageable memory through the threaded stager.
    const float *dev_map = nullptr;
    cudaPointerAttributes attr;
    memset(&attr, 0, sizeof(attr));
    if (cudaPointerGetAttributes(&attr, pafmap) != cudaSuccess) {
        cudaGetLastError();
        attr.type = cudaMemoryTypeUnregistered;
    }
    if (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged) {
        if (attr.type == cudaMemoryTypeDevice && attr.device != h->device)
            return fail(IBP_ERR_ARG, "process_paf: the limb map lives on device %d, the handle on device %d (IBP_DEVICE)", attr.device, h->device);
        dev_map = pafmap;
    } else if (n > 0 && legacy_sample_over_bus(attr, counts, h->P.max_steps, f1, f2, f3, y_lo, y_hi)) {
        // pinned / registered memory and few pairs: the kernel reads its samples (one 32-byte sector each) straight out
        // of the caller's array over the bus; nothing is copied
        dev_map = (const float *)attr.devicePointer;
    } else if (n > 0) {
        const size_t row_bytes = (size_t)f2 * f3 * sizeof(float);
        rc = ensure_stage(h, (size_t)f1 * row_bytes, st);
        if (rc != IBP_OK) return rc;
        const int r0 = std::max(0, std::min(y_lo, f1 - 1) - 1), r1 = std::min(f1 - 1, std::max(y_hi, 0) + 1);
        const size_t off = (size_t)r0 * row_bytes, bytes = (size_t)(r1 - r0 + 1) * row_bytes;
        char *dst = (char *)h->d_stage + off;
        const char *src = (const char *)pafmap + off;
        static const bool no_stager = getenv("IBP_NO_STAGER") != nullptr;
        bool sent = false;
        if (attr.type == cudaMemoryTypeHost) {  // pinned / registered: the copy engine reads it directly
            CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
            sent = true;
        } else if (bytes >= 4 * HostStager::kChunk && !no_stager) {
            HostStager *stager = HostStager::get(h->device);
            // the stager's streams do not order after `st`: the zero fill of a fresh staging buffer must be over first
            if (stager && cudaStreamSynchronize(st) == cudaSuccess) sent = stager->upload(src, dst, bytes, st);
            if (!sent) cudaGetLastError();
        }
        if (!sent) CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
        dev_map = (const float *)h->d_stage;
    } else {
        rc = ensure_stage(h, 256, st);
        if (rc != IBP_OK) return rc;
        dev_map = (const float *)h->d_stage;  // no peaks: nothing is sampled
    }
    rc = ibp_assign(h, dev_map, IBP_LIMB_FULLRES_HWC, 1, f1, f2, f3, min_img_size, st);
    if (rc != IBP_OK) return rc;
    // results: humans + status only; the peak list is already on the host (ibp_set_peaks filled h_line)
    CU(cudaMemcpyAsync(h->h_n_humans, h->d_n_humans, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_humans, h->d_humans, (size_t)h->L.max_humans * sizeof(ibp_human), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h->h_status, h->d_status, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    g_view.batch = 1;
    g_view.peak_stride = h->line_stride;
    g_view.human_stride = h->L.max_humans;
    g_view.n_peaks = h->h_n_peaks;
    g_view.part_count = h->h_part_count;
    g_view.peaks = h->h_line;
    g_view.n_humans = h->h_n_humans;
    g_view.humans = h->h_humans;
    g_view.status = h->h_status;
    g_view_valid = true;
    g_legacy_status = h->h_status[0];
    if (g_legacy_status & (IBP_ST_PEAK_OVERFLOW | IBP_ST_CAND_OVERFLOW | IBP_ST_BODY_OVERFLOW | IBP_ST_HUMAN_OVERFLOW))
        return fail(IBP_ERR_CAPACITY, "process_paf: device buffer overflow (status 0x%x)", g_legacy_status);
    return IBP_OK;  // the reference returns 0 (pafprocess.cpp:284)
}

int ibp_get_num_humans(void) { return g_view_valid ? g_view.n_humans[0] : 0; }
int ibp_get_part_peak_id(int skeleton_id, int part_id) {
    if (!g_view_valid || skeleton_id < 0 || skeleton_id >= g_view.n_humans[0] || part_id < 0 || part_id >= kParts) return -1;
    return g_view.humans[skeleton_id].peak_id[part_id];
}
float ibp_get_score(int skeleton_id) {
    if (!g_view_valid || skeleton_id < 0 || skeleton_id >= g_view.n_humans[0]) return 0.0f;
    return g_view.humans[skeleton_id].score;
}
int ibp_get_part_x(int cid) { return (g_view_valid && cid >= 0 && cid < g_view.n_peaks[0]) ? g_view.peaks[cid].x : 0; }
int ibp_get_part_y(int cid) { return (g_view_valid && cid >= 0 && cid < g_view.n_peaks[0]) ? g_view.peaks[cid].y : 0; }
float ibp_get_part_score(int cid) { return (g_view_valid && cid >= 0 && cid < g_view.n_peaks[0]) ? g_view.peaks[cid].score : 0.0f; }
uint32_t ibp_get_status(void) { return g_legacy_status; }

// ------------------------------------------------------------------------------------------------ allocation-free peaks
// No counterpart in the reference (include/ibp.h §5).  Nothing in here
// allocates: no cudaMalloc / cudaFree / cudaMallocHost / cudaStreamCreate / cudaEventCreate, no std::vector.
namespace {
constexpr int kNoAllocMaxDevices = 64;
NoAllocScratch *g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};  // DEVICE address of g_noalloc_scratch, per device
std::mutex g_noalloc_mu;
}  // namespace

int ibp_find_peaks_noalloc(const float *maps, int chan_stride, int first_chan, int batch, int fh, int fw, ibp_peak *out_peaks,
                           int32_t *out_part_count, int32_t *out_n_peaks, uint32_t *out_status, void *stream) {
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(IBP_ERR_ARG, "ibp_find_peaks_noalloc: device %d not supported (max %d)", dev, kNoAllocMaxDevices);
    auto kern = peaks_kernel<float, IBP_PEAKS_CROSS_REFINE>;
    {
        // set-up, once per device, with that device current: the DEVICE address of the scratch array (its name is only a
        // host-side stand-in here) and the kernel's shared-memory opt-in
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[dev]) {
            void *p = nullptr;
            CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            if (!p) return fail(IBP_ERR_CUDA, "ibp_find_peaks_noalloc: cudaGetSymbolAddress returned a null address");
            CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
            g_noalloc_ptr[dev] = static_cast<NoAllocScratch *>(p);
        }
    }
    if (batch == 0) return IBP_OK;  // set-up only (callers about to capture a graph)
    if (!maps || !out_peaks || !out_part_count || !out_n_peaks || !out_status) return fail(IBP_ERR_ARG, "ibp_find_peaks_noalloc: null argument");
    if (batch < 0 || batch > IBP_NOALLOC_MAX_BATCH)
        return fail(IBP_ERR_ARG, "ibp_find_peaks_noalloc: batch %d outside [0,%d] (IBP_NOALLOC_MAX_BATCH: the static scratch array is sized for it)",
                    batch, IBP_NOALLOC_MAX_BATCH);
    if (fh < 1 || fw < 1 || fw > 2048 || fh > 32767 || first_chan < 0 || first_chan + kParts > chan_stride)
        return fail(IBP_ERR_SHAPE, "ibp_find_peaks_noalloc: bad shape (chan_stride %d, first_chan %d, %dx%d; fw <= 2048)", chan_stride, first_chan, fh, fw);
    NoAllocScratch *S = g_noalloc_ptr[dev];
    cudaStream_t st = (cudaStream_t)stream;

    // the geometry of launch_peaks' one-CTA-per-band route, default parameters (ibp_default_params)
    PeakArgs a;
    memset(&a, 0, sizeof(a));
    a.src = maps;
    a.dtype = IBP_F32;
    a.image_stride = (size_t)chan_stride * fh * fw;
    a.chan0 = first_chan;
    a.twin_base = -1;
    a.h = fh;
    a.w = fw;
    a.halo = 2;
    const int rows_target = std::max<int>(4 * a.halo, (int)(kPeakBandTarget / ((size_t)fw * 4)));
    int nbands = (fh + rows_target - 1) / rows_target;
    a.rows_per_band = (fh + nbands - 1) / nbands;
    a.nbands = (fh + a.rows_per_band - 1) / a.rows_per_band;
    if (a.nbands > IBP_NOALLOC_MAX_BANDS)
        return fail(IBP_ERR_SHAPE, "ibp_find_peaks_noalloc: a %dx%d map needs %d bands of %d rows, the static scratch array holds %d "
                                   "(IBP_NOALLOC_MAX_BANDS); use ibp_find_peaks", fh, fw, a.nbands, a.rows_per_band, IBP_NOALLOC_MAX_BANDS);
    const int rows_buf = std::min(a.rows_per_band + 2 * a.halo, fh);
    size_t band_bytes = std::max((size_t)rows_buf * fw * 4, (size_t)(kParts * a.nbands + 1) * 4);
    band_bytes = std::max(band_bytes, (size_t)batch * 4);
    band_bytes = (band_bytes + 15) & ~(size_t)15;
    a.band_bytes = (int)band_bytes;
    a.ring_bytes = 0;
    a.up = 4;
    a.win = 2;
    a.thr = 0.1f;
    a.cap = IBP_NOALLOC_CAP;
    a.flavour = IBP_PEAKS_CROSS_REFINE;
    a.raw = 0;
    a.band_peaks = S->band_peaks;
    a.band_subpix = nullptr;
    a.band_count = S->band_count;
    a.ticket = S->ticket;
    a.peaks = S->peaks;
    a.subpix = nullptr;
    a.part_count = out_part_count;
    a.line = out_peaks;
    a.line_stride = kParts * IBP_NOALLOC_CAP;
    a.n_peaks = out_n_peaks;
    a.status = out_status;
    a.batch_ticket = &S->batch_ticket;
    a.timeline = nullptr;
    a.order = S->order;
    a.batch = batch;
    const int side = 2 * a.win + 1;
    size_t scratch = (((size_t)(kPeakThreads / 32) * side * side * a.up * 4) + 7) & ~(size_t)7;
    scratch += (size_t)a.cap * 8;
    scratch = (scratch + 15) & ~(size_t)15;
    const size_t smem = band_bytes + scratch;
    if (smem > kMaxDynSmem) return fail(IBP_ERR_SHAPE, "ibp_find_peaks_noalloc: rows of %d floats do not fit shared memory", fw);

    // the array keeps the last call's contents (a call that died mid-way leaves tickets behind): clear what must start at
    // zero, on the caller's stream; the status words are the caller's
    const size_t tail = sizeof(NoAllocScratch) - offsetof(NoAllocScratch, band_count);
    CU(cudaMemsetAsync(S->band_count, 0, tail, st));
    CU(cudaMemsetAsync(out_status, 0, (size_t)batch * sizeof(uint32_t), st));
    kern<<<dim3(a.nbands, kParts, batch), kPeakThreads, smem, st>>>(a);
    CU(cudaGetLastError());
    return IBP_OK;
}

}  // extern "C"
