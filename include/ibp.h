/* ibp.h — C ABI of the B200-native keypoint-assignment path (network output -> skeletons).
 *
 * One shared library, libibp_b200.so, hand-written sm_100a CUDA behind plain-C entry points: pointers and sizes
 * only, no torch / numpy / C++ types.  Each entry point names the reference interface it replaces
 * (paths relative to the reference repository sokunmin/Improved-Body-Parts).
 *
 * Three groups:
 *   1. the seven legacy calls of utils/pafprocess/pafprocess.h:70-76 (what SWIG binds today) — host pointers,
 *      process-global result state, same argument meaning, same getter protocol;
 *   2. stage entry points on DEVICE pointers (fuse, peaks, limb up-sampling, assignment) for callers that keep
 *      tensors resident — what utils/parse_skeletons.py's predict_refactor / heatmap_nms and evaluate.py:77-80
 *      compute on the host today;
 *   3. the batched pipeline (net output of B images in, packed skeletons out) that bench.py measures.
 *
 * There is no CPU fallback: every compute call returns IBP_ERR_NO_DEVICE / IBP_ERR_CUDA when no sm_100 device
 * can run it.  Nothing aborts; ibp_last_error() gives the message of the last failing call on this thread.
 */
#ifndef IBP_B200_H
#define IBP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IBP_NUM_PARTS 18  /* pafprocess.h:11 NUM_PART; config/config.py:64 */
#define IBP_NUM_LIMBS 30  /* pafprocess.h:19 COCOPAIRS_SIZE; config/config.py:125 */
#define IBP_NUM_LAYERS 50 /* config/config.py:128: [0,30) limb, [30,48) keypoint, [48,50) background */

/* return codes (the reference always returns 0 and has undefined behaviour on bad input) */
#define IBP_OK 0
#define IBP_ERR_SHAPE (-1)     /* p3 < 5, f3 < 30, non-positive sizes */
#define IBP_ERR_PART (-2)      /* part id outside [0,18) in the joint list */
#define IBP_ERR_CAPACITY (-3)  /* a fixed-capacity device buffer overflowed; see ibp_limits and status bits */
#define IBP_ERR_CUDA (-4)      /* CUDA runtime error, message in ibp_last_error() */
#define IBP_ERR_ARG (-5)       /* null pointer / bad enum / batch larger than the handle's limit */
#define IBP_ERR_NO_DEVICE (-6) /* no CUDA device: this library has no CPU path */

/* element type of the network output handed to the fuse / pipeline calls */
#define IBP_F32 0
#define IBP_F16 1 /* apex amp O1 output, utils/parse_skeletons.py:75 */

/* where the limb values come from in ibp_assign */
#define IBP_LIMB_FULLRES_HWC 0 /* (H,W,30) float32 per image, the reference layout (pafprocess.cpp:9 PAF macro) */
#define IBP_LIMB_FEAT_PLANAR 1 /* (C,h,w) planar feature-resolution maps, bicubic x`up` evaluated at the samples */

/* peak extraction flavours */
#define IBP_PEAKS_CROSS_REFINE 0 /* heatmap_nms: 4-connected NMS, strict > thr, 5x5 bicubic patch refine
                                    (utils/parse_skeletons.py:106-176) */
#define IBP_PEAKS_BOX_CENTROID 1 /* find_peaks: 3x3 NMS, >= thr, 5x5 centroid refine
                                    (utils/parse_skeletons.py:286-321, utils/util.py:177-213) */

#define IBP_PEAKS_CROSS_RAW 2    /* heatmap_nms(bool_refine_center=False): the cross NMS of flavour 0 without the patch
                                    refine — x, y = the feature-map pixel, score = the map value there
                                    (utils/parse_skeletons.py:163-166); the caller maps to (x + 0.5) * up - 0.5 (l.122-123) */

/* per-image status bits (ibp_batch_view.status) */
#define IBP_ST_PEAK_OVERFLOW 0x1u    /* more peaks of one part than limits.max_peaks_per_part */
#define IBP_ST_CAND_OVERFLOW 0x2u    /* more accepted pairs on one limb than limits.max_candidates_per_limb */
#define IBP_ST_BODY_OVERFLOW 0x4u    /* more simultaneous partial skeletons than limits.max_bodies */
#define IBP_ST_HUMAN_OVERFLOW 0x8u   /* more humans than limits.max_humans */
#define IBP_ST_TIE_EMULATED 0x10u    /* a limb had > 16 candidates with duplicate sort keys; order taken from the
                                        libstdc++ introsort emulation (pafprocess.cpp:109) */
#define IBP_ST_STALLED 0x40u         /* internal: the grouping of an image never saw all 30 limbs finish (ibp_fetch fails) */
#define IBP_ST_TIE_UNDEFINED 0x20u   /* ...and the reference's std::sort would have read out of bounds there */

/* Constants the reference compiles in (pafprocess.h:6-18) or reads from utils/config:16-27. */
typedef struct ibp_params {
    float limb_sample_thresh; /* THRESH_PAF_SCORE 0.1        | thre2 */
    float step_ratio;         /* THRESH_PAF_STEP_RATIO 0.8   | connect_ration */
    int32_t min_parts;        /* THRESH_PART_CNT 2 */
    float min_human_score;    /* THRESH_SKELETON_SCORE 0.45 */
    int32_t max_steps;        /* STEP_PAF 20                 | mid_num */
    int32_t len_rate;         /* LIMB_LENGTH_RATE 16         | len_rate */
    float merge_tolerance;    /* MIN_SCORE_TOLERANCE 0.7     | connection_tole */
    float weight_limb;        /* PAF_OUT_WEIGHTS[0] 0.5 */
    float weight_a;           /* PAF_OUT_WEIGHTS[1] 0.25 */
    float weight_b;           /* PAF_OUT_WEIGHTS[2] 0.25 */
    float peak_thresh;        /* 0.1: literal at parse_skeletons.py:139 | thre1 for the box/centroid flavour */
    int32_t upsample;         /* 4: model_cfg['stride'] (evaluate.py:76-80) */
    int32_t refine_window;    /* 2: win_size at parse_skeletons.py:135 | offset_radius */
} ibp_params;

/* Fixed capacities of the device-side variable-length outputs.  Defaults (ibp_default_limits): batch 64, 128 peaks per
 * part, 1024 candidates per limb, 128 bodies, 128 humans, 128x128 maps.  With float32 network output, <= 128 peaks per
 * part and grouping tables that fit behind a limb plane in shared memory (128 bodies at 128x128) ibp_postprocess_batch
 * runs the persistent streaming kernels; larger limits (crowds) take the one-CTA-per-item kernels.  Same results. */
typedef struct ibp_limits {
    int32_t max_batch;               /* images per call */
    int32_t max_peaks_per_part;      /* per image and joint type */
    int32_t max_candidates_per_limb; /* accepted pairs kept for the sort, power of two */
    int32_t max_bodies;              /* live partial skeletons during grouping */
    int32_t max_humans;              /* humans reported per image */
    int32_t max_feat_h, max_feat_w;  /* largest feature map handed to the pipeline */
} ibp_limits;

typedef struct ibp_peak { /* one row of the joint list (evaluate.py:99-103) after pafprocess.cpp:33-38 */
    int32_t x, y;         /* truncated to int like the reference's Peak (pafprocess.h:29-34) */
    float score;
    int32_t part;
} ibp_peak;

typedef struct ibp_human { /* one row of the reference's skeleton table (pafprocess.cpp:16, pafprocess.h:36-45) */
    int32_t peak_id[IBP_NUM_PARTS]; /* get_part_peak_id: id or -1 */
    float score;                    /* get_score: total / count */
    float total;                    /* slot 18 .score */
    int32_t count;                  /* slot 19 .count */
    float reach;                    /* slot 19 .length */
} ibp_human;

typedef struct ibp_batch_view { /* host pointers into pinned memory owned by the handle, valid until the next call */
    int32_t batch;
    int32_t peak_stride;        /* peaks of image b start at peaks + b*peak_stride */
    int32_t human_stride;       /* humans of image b start at humans + b*human_stride */
    const int32_t *n_peaks;     /* [batch] */
    const int32_t *part_count;  /* [batch][18] peaks per joint type; peaks are stored part-major ("line" order) */
    const ibp_peak *peaks;      /* [batch][peak_stride], index == peak id */
    const int32_t *n_humans;    /* [batch] */
    const ibp_human *humans;    /* [batch][human_stride] */
    const uint32_t *status;     /* [batch] IBP_ST_* bits */
} ibp_batch_view;

typedef struct ibp_handle ibp_handle;

/* ---------------------------------------------------------------------------------------------------------------
 * 1. Legacy protocol — replaces utils/pafprocess/pafprocess.h:70-76 as bound by utils/pafprocess/pafprocess.i:14.
 *    HOST pointers.  Results live in a process-global handle until the next ibp_process_paf (the reference keeps
 *    them in two globals, pafprocess.cpp:16-17); not thread-safe, exactly like the reference.
 * ------------------------------------------------------------------------------------------------------------- */
/* pafprocess.h:70 process_paf: peaks (p1,p2,p3>=5) rows [x,y,score,id,part]; pafmap (f1=H,f2=W,f3>=30) HWC. */
int ibp_process_paf(int p1, int p2, int p3, const float *peaks, int f1, int f2, int f3, const float *pafmap,
                    int min_img_size);
int ibp_get_num_humans(void);                            /* pafprocess.h:71 */
int ibp_get_part_peak_id(int skeleton_id, int part_id);  /* pafprocess.h:72 */
float ibp_get_score(int skeleton_id);                    /* pafprocess.h:73 */
int ibp_get_part_x(int cid);                             /* pafprocess.h:74 */
int ibp_get_part_y(int cid);                             /* pafprocess.h:75 */
float ibp_get_part_score(int cid);                       /* pafprocess.h:76 */
uint32_t ibp_get_status(void);                           /* IBP_ST_* bits of the last legacy call (no counterpart) */

/* ---------------------------------------------------------------------------------------------------------------
 * Handles, defaults, errors.
 * ------------------------------------------------------------------------------------------------------------- */
void ibp_default_params(ibp_params *p);
void ibp_default_limits(ibp_limits *l);
int ibp_create(ibp_handle **out, int device, const ibp_params *params /*NULL=defaults*/,
               const ibp_limits *limits /*NULL=defaults*/);
void ibp_destroy(ibp_handle *h);
const char *ibp_last_error(void);
int ibp_device_count(void); /* 0 when there is no usable CUDA device */

/* ---------------------------------------------------------------------------------------------------------------
 * 2. Stage entry points.  All pointers are DEVICE pointers on the handle's device unless named host_*; `stream`
 *    is a cudaStream_t passed as void* (NULL = default stream).  Calls are asynchronous on that stream.
 * ------------------------------------------------------------------------------------------------------------- */
/* K1 — flip/channel-permutation average: utils/parse_skeletons.py:80-96 (predict_refactor) and :232-248 (predict).
 * net_out (batch,2,50,h,w) of `dtype`; fused (batch,50,h,w) float32 planar:
 *   fused[c] = (net_out[0][c] + mirror_x(net_out[1][perm[c]])) / 2, averaged in `dtype` like numpy does.
 * flip_avg == 0 copies net_out[:,0] (parse_skeletons.py:94-96). */
int ibp_fuse_flip(ibp_handle *h, const void *net_out, int dtype, int batch, int fh, int fw, int flip_avg,
                  float *fused, void *stream);

/* K1, multi-scale — scale-search / flip fusion at image resolution: utils/parse_skeletons.py:180-283 (predict), what
 * follows the model calls.  One entry per scale of the search (l.186, l.194): net_out (batch,2,50,fh,fw) float32 is the
 * network output for the input scaled to (sh,sw) and padded to (fh*4, fw*4) (l.204-208).  Per scale the maps are flip-
 * averaged, resized x4 (cv2 INTER_CUBIC), cropped to (sh,sw), resized to (img_h,img_w) and added / n_scales (l.232-281).
 * out (batch,50,img_h,img_w) planar float64: channels [0,30) = paf_avg, [30,50) = heatmap_avg.  heat32 (optional, may be
 * NULL): (batch,18,img_h,img_w) float32 copy of the keypoint planes, what find_peaks (l.291) works on. */
typedef struct ibp_scale_input {
    const float *net_out;
    int32_t fh, fw;
    int32_t sh, sw;
} ibp_scale_input;
int ibp_fuse_multiscale(ibp_handle *h, const ibp_scale_input *scales, int n_scales, int batch, int img_h, int img_w,
                        int flip_avg, double *out, float *heat32, void *stream);

/* Second operator mode (the operators BASELINE.json's north_star names where the reference's live
 * path uses others).  NOT what utils/parse_skeletons.py computes: the same scale-search / flip fusion as above
 * (l.180-283) with cv2.INTER_LINEAR in place of INTER_CUBIC in both resizes (l.252-263, l.276-277), the two bilinear
 * resizes composed into one <= 3 x 3 tap operator on the feature maps.  Same arguments and outputs as
 * ibp_fuse_multiscale; feature maps of at least 3 x 3.  Checked against its own CPU oracle (oracle/np_oracle.py:
 * predict_fuse_linear, which is held to cv2.resize) to 1e-4 relative, not bit for bit. */
int ibp_fuse_multiscale_linear(ibp_handle *h, const ibp_scale_input *scales, int n_scales, int batch, int img_h, int img_w,
                               int flip_avg, double *out, float *heat32, void *stream);

/* Second operator mode, smoothing: utils/util.py:103-174 (GaussianSmoothing, unused by the reference's live path) applied
 * the way its docstring says, to F.pad(x, (p,p,p,p), mode='reflect') with p = ksize / 2: separable Gaussian of `planes`
 * planar (img_h,img_w) float32 maps, odd ksize <= 15.  The smoothed maps then go through ibp_find_peaks (flavour 1:
 * 3x3 max-suppression into per-channel candidate lists).  Oracle: oracle/np_oracle.py:gaussian_smooth, pinned by the
 * reference class's own weights (tests/golden/modes_12x.npz); tolerance 1e-5 absolute. */
int ibp_smooth_gaussian(ibp_handle *h, const float *maps, float *out, int planes, int img_h, int img_w, int ksize, float sigma,
                        void *stream);

/* A4 — limb-map up-sampling: evaluate.py:77-80 / demo_image.py:94-97, cv2.resize(pafs, fx=up, fy=up, INTER_CUBIC).
 * fused (batch,chan_stride,h,w) planar, limb channels first; out (batch, h*up, w*up, 30) HWC float32. */
int ibp_upsample_limbs(ibp_handle *h, const float *fused, int chan_stride, int batch, int fh, int fw, int up,
                       float *out_hwc, void *stream);

/* K2 — peaks: utils/parse_skeletons.py:126-176 heatmap_nms (flavour 0) or :286-321 find_peaks (flavour 1).
 * maps (batch,chan_stride,h,w) planar float32, keypoint channel j at index first_chan + j.  Peaks are left in the
 * handle's device buffers (part-major, row-major inside a part, ids = running counter, parse_skeletons.py:172-173)
 * for ibp_assign / ibp_fetch. */
int ibp_find_peaks(ibp_handle *h, const float *maps, int chan_stride, int first_chan, int batch, int fh, int fw,
                   int flavour, void *stream);

/* Load an explicit joint list instead (legacy path, tests): host_peaks (n,5) rows [x,y,score,id,part] for ONE
 * image slot `image`; ids are arrival order (pafprocess.cpp:34).  Synchronous host->device copy. */
int ibp_set_peaks(ibp_handle *h, int image, const float *host_peaks, int n, int row_stride);

/* K3+K4 — pair scoring, ordering, greedy connection, grouping, filter: utils/pafprocess/pafprocess.cpp:52-285.
 * limb_src layout per `mode`; for IBP_LIMB_FULLRES_HWC: (batch,H,W,C>=30) with H=dim0, W=dim1, C=dim2;
 * for IBP_LIMB_FEAT_PLANAR: (batch,dim2,dim0,dim1) planar with `up` from the handle's params.
 * min_img_size: the reference passes the image height (evaluate.py:110). */
int ibp_assign(ibp_handle *h, const float *limb_src, int mode, int batch, int dim0, int dim1, int dim2,
               int min_img_size, void *stream);

/* ---------------------------------------------------------------------------------------------------------------
 * 3. Batched pipeline: everything evaluate.py:75-129 does after the model call, for `batch` images.
 *    net_out (batch,2,50,h,w); on_host != 0 means a HOST pointer.  Pageable: the 48 limb + keypoint planes of each half are
 *    copied in-stream (the two background planes are never read on this path).  Pinned (page-locked, as cudaHostAlloc /
 *    torch pin_memory give it): only the 18 keypoint planes of each half are copied; the limb kernel reads the limb rows it
 *    needs straight out of the caller's buffer over the bus (about 36 % fewer bytes), so that buffer must stay unchanged
 *    until ibp_fetch / ibp_fetch_coco has returned for this batch (IBP_NO_ZEROCOPY=1 in the environment forces the copy).
 *    ibp_fetch waits for the stream, copies the packed results to the handle's pinned buffers and fills `view`.
 * ------------------------------------------------------------------------------------------------------------- */
int ibp_postprocess_batch(ibp_handle *h, const void *net_out, int on_host, int dtype, int batch, int fh, int fw,
                          int min_img_size, void *stream);
int ibp_fetch(ibp_handle *h, int batch, ibp_batch_view *view, void *stream);
/* Batched result materialisation — evaluate.py:182-209 append_result on the device: every
 * human of the batch as one row of IBP_COCO_ROW floats, 17 x (x, y, v) in COCO order (ORDER_COCO, evaluate.py:40; v = 1
 * and the peak's integer position for a found part, zeros otherwise) followed by the human's score (get_score).  Rows
 * of all images are compacted on the device; the call copies offsets[batch+1] and exactly offsets[batch] rows.
 * Valid after ibp_postprocess_batch / ibp_assign; host pointers into pinned memory owned by the handle. */
#define IBP_COCO_ROW 52
typedef struct ibp_coco_view {
    int32_t batch;
    const int32_t *offsets; /* [batch + 1]: rows of image b are rows[offsets[b] .. offsets[b+1]) */
    const float *rows;      /* [offsets[batch]][IBP_COCO_ROW] */
    const uint32_t *status; /* [batch] IBP_ST_* bits */
} ibp_coco_view;
int ibp_fetch_coco(ibp_handle *h, int batch, ibp_coco_view *view, void *stream);

/* Image preparation — utils/parse_skeletons.py:52-72 after the cv2.resize: pad right/down to a
 * multiple of `max_downsample` with `pad_value` (utils/util.py:44-65 padRightDownCorner), np.float32(img / 255), and the
 * horizontally flipped twin.  DEVICE pointers: img (batch,h,w,3) uint8 HWC; out (batch, flip ? 2 : 1, hp, wp, 3) float32
 * HWC with hp, wp = h, w rounded up to the multiple (a dimension of 1 stays 1: util.py:59-62 tiles an empty slice there)
 * — the (2*batch, hp, wp, 3) tensor the model is called on (l.72). */
int ibp_prepare_image(ibp_handle *h, const uint8_t *img, int batch, int img_h, int img_w, int max_downsample,
                      int pad_value, int flip, float *out, void *stream);

/* Image resize of the scale search — utils/parse_skeletons.py:204 (predict) / :52 (predict_refactor):
 * cv2.resize(image, (0, 0), fx=fx, fy=fy, interpolation=cv2.INTER_CUBIC) on uint8 HWC images, OpenCV's own 8-bit path
 * (11-bit fixed-point taps, integer horizontal pass, float vertical pass, integer tail; csrc/k_resize.cuh), bit-exact
 * against cv2 with IPP off.  DEVICE pointers: img (batch,h,w,3), out (batch,out_h,out_w,3) with (out_h, out_w) =
 * ibp_resized_size (rint(h*fy), rint(w*fx), cv2's saturate_cast<int>); any other size is IBP_ERR_SHAPE. */
void ibp_resized_size(int img_h, int img_w, double fx, double fy, int *out_h, int *out_w);
int ibp_resize_image(ibp_handle *h, const uint8_t *img, int batch, int img_h, int img_w, double fx, double fy, uint8_t *out,
                     int out_h, int out_w, void *stream);

/* The rotation_search branch (knob utils/config:15) — utils/parse_skeletons.py:215-218, :265-267
 * (predict) and :63-66, :98-100 (predict_refactor).
 * ibp_rotation_matrix: cv2.getRotationMatrix2D(center=(cx, cy), angle [degrees], scale) -> m[6] (2x3, row-major, double).
 * ibp_warp_affine: cv2.warpAffine(src, M, (0, 0)) with the defaults (INTER_LINEAR, BORDER_CONSTANT 0, output size = source
 *   size) for float32 DEVICE data of `channels` channels addressed by element strides (batch, channel, row, pixel), the
 *   same for src and dst: planar maps (B,C,h,w) = strides (C*h*w, h*w, w, 1); an HWC image (B,h,w,3) = (h*w*3, 1, w*3, 3).
 *   m = the FORWARD matrix as the caller would pass it to cv2 (the library inverts it like cv::warpAffine does).
 * ibp_crop_resize_accumulate: one (scale, rotation) item of predict's tail when a warp sits between the two resizes
 *   (l.272-281): maps (batch,50,ph,pw) planar float32 at the padded input size are cropped to (sh,sw), resized to
 *   (img_h,img_w) with cv2's INTER_CUBIC, divided by `n_items` in float32 and ADDED to out (batch,50,img_h,img_w) float64. */
void ibp_rotation_matrix(double cx, double cy, double angle_deg, double scale, double *m6);
int ibp_warp_affine(ibp_handle *h, const float *src, float *dst, int batch, int channels, int hh, int ww, int64_t batch_stride,
                    int64_t chan_stride, int64_t row_stride, int64_t pix_stride, const double *m6, void *stream);
int ibp_crop_resize_accumulate(ibp_handle *h, const float *maps, int batch, int ph, int pw, int sh, int sw, int img_h, int img_w,
                               int n_items, double *out, void *stream);

/* Sub-pixel coordinates of the IBP_PEAKS_BOX_CENTROID flavour (utils/util.py:188-213 returns float64 x, y): after
 * ibp_fetch, copies image `image`'s (x, y) pairs in peak-id order into host_xy[capacity][2].  Returns the number of
 * peaks written (>= 0) or a negative error code. */
int ibp_fetch_subpix(ibp_handle *h, int image, double *host_xy, int capacity, void *stream);
/* ---------------------------------------------------------------------------------------------------------------
 * 4. The reference's PYTHON assignment route in float64 — utils/parse_skeletons.py:324-410 (find_connections) and
 *    :413-599 (find_humans), what evaluate.py:88-90 runs without --run_cpp and the only grouping of the original
 *    route (predict + find_peaks).  Peaks come from ibp_find_peaks(IBP_PEAKS_BOX_CENTROID) (device resident) or from
 *    ibp_set_peaks64.  Parameters are the Python floats of utils/config (the C++ constants are float32).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct ibp_py_params {
    double thre2;           /* 0.1  utils/config:17 */
    double connect_ration;  /* 0.8  utils/config:18 */
    int32_t mid_num;        /* 20   utils/config:20 */
    int32_t len_rate;       /* 16   utils/config:21 */
    double connection_tole; /* 0.7  utils/config:22 */
    double min_score;       /* 0.45 literal at parse_skeletons.py:594 */
    int32_t min_parts;      /* 2    literal at parse_skeletons.py:594 */
} ibp_py_params;
void ibp_default_py_params(ibp_py_params *p);
/* all_peaks of ONE image as host rows (n,4) [x, y, score, id] float64 with their joint type part[n], in list order. */
int ibp_set_peaks64(ibp_handle *h, int image, const double *rows, const int32_t *part, int n);
/* find_connections: paf (batch, chan_stride>=30, img_h, img_w) planar float64 DEVICE pointer (ibp_fuse_multiscale's
 * `out`); img_height is the reference's third argument (evaluate.py:89). */
int ibp_find_connections(ibp_handle *h, const ibp_py_params *params, const double *paf, int chan_stride, int batch,
                         int img_h, int img_w, int img_height, void *stream);
/* connected_limbs of one image: rows (30, max_peaks_per_part, 6) [src_id, dst_id, score, i, j, length], counts[30]. */
int ibp_fetch_connections(ibp_handle *h, int image, double *rows, int32_t *counts, void *stream);
/* explicit connected_limbs + joint_candidates[:, 2] for find_humans (same layout as ibp_fetch_connections). */
int ibp_set_connections(ibp_handle *h, int image, const double *rows, const int32_t *counts, const double *cand_score,
                        int n_cand);
int ibp_find_humans(ibp_handle *h, const ibp_py_params *params, int batch, void *stream);
/* person_to_joint_assoc of one image: people (n,20,2) float64; returns n (>= 0) or a negative error code. */
int ibp_fetch_people(ibp_handle *h, int image, double *people, int capacity, void *stream);

/* ---------------------------------------------------------------------------------------------------------------
 * 5. Allocation-free peak extraction.  NO COUNTERPART IN THE REFERENCE: added for
 *    callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *    Same computation as ibp_find_peaks with IBP_PEAKS_CROSS_REFINE (heatmap_nms, 4-connected NMS > thresh, 5x5 bicubic x4
 *    refine, running ids) with the default parameters (thresh 0.1, up-sampling 4, window 2), but without a handle: the
 *    call allocates nothing — no cudaMalloc / cudaFree / cudaMallocHost, no allocating library routine.  Its scratch (the
 *    per-band peak lists, the per-part lists, band counts and tickets) lives in ONE file-scope `static __device__` array
 *    of IBP_NOALLOC_SCRATCH_BYTES (about 1.3 MB) that every GPU loading the library holds until the process exits.
 *    Limits (rejected with IBP_ERR_ARG / IBP_ERR_SHAPE and a message in ibp_last_error() beyond them):
 *      batch <= IBP_NOALLOC_MAX_BATCH (8); at most IBP_NOALLOC_MAX_BANDS (8) bands per plane, i.e.
 *      fh <= 8 * max(8, 4096 / fw) rows (256 rows at fw = 128); fw <= 2048; IBP_NOALLOC_CAP (64) peaks per joint type.
 *    All pointers are DEVICE pointers on the calling thread's CURRENT device:
 *      maps (batch, chan_stride, fh, fw) planar float32, keypoint j in channel first_chan + j           [in]
 *      out_peaks (batch, 18 * IBP_NOALLOC_CAP) rows, part-major, index == peak id                      [out]
 *      out_part_count (batch, 18), out_n_peaks (batch), out_status (batch) IBP_ST_* bits               [out]
 *    Asynchronous on `stream`.  The first call on a device does the one-time set-up (address of the scratch array,
 *    kernel attributes; nothing is allocated) — under graph capture make one call with batch == 0 beforehand, which
 *    does only that.  EVERY call shares the one scratch array: calls that overlap on the GPU (different streams,
 *    different threads, concurrent graph launches) RACE; serialise them.  The array keeps the last call's contents;
 *    the call clears what must start at zero itself (memset nodes on `stream`). */
#define IBP_NOALLOC_MAX_BATCH 8
#define IBP_NOALLOC_MAX_BANDS 8
#define IBP_NOALLOC_CAP 64
#define IBP_NOALLOC_SCRATCH_BYTES (1360 * 1024)
int ibp_find_peaks_noalloc(const float *maps, int chan_stride, int first_chan, int batch, int fh, int fw,
                           ibp_peak *out_peaks, int32_t *out_part_count, int32_t *out_n_peaks, uint32_t *out_status,
                           void *stream);

/* number of kernels this library has launched through `h` since creation (bench.py's gpu_launches) */
int64_t ibp_launch_count(const ibp_handle *h);


/* Per-kernel device times, measured with CUDA events recorded on the launching stream around each kernel of the
 * last call(s).  Stage indices below; ms[i] < 0 when the stage did not run.  ibp_get_stage_ms waits for the events. */
#define IBP_STAGE_FUSE 0
#define IBP_STAGE_PEAKS 1
#define IBP_STAGE_PACK 2
#define IBP_STAGE_LIMBS 3
#define IBP_STAGE_GROUP 4
#define IBP_STAGE_UPSAMPLE 5
#define IBP_STAGE_SMOOTH 6
#define IBP_NUM_STAGES 7
int ibp_set_profiling(ibp_handle *h, int on);
int ibp_get_stage_ms(ibp_handle *h, float *ms /* [IBP_NUM_STAGES] */);

#ifdef __cplusplus
}
#endif
#endif /* IBP_B200_H */
