/*
 * rtc_b200.h -- C ABI of librtc_b200.so, the B200-native render loop.
 *
 * The reference has no plugin / FFI interface: its render step is the inline
 * pthread fan-out of main() (reference main.c:285-302, 12 x rowThread,
 * main.c:81-104).  render_cuda() is the drop-in for exactly those lines: same
 * inputs (Scene by value -> pointer, the camera basis of main.c:252-255, the
 * Triangle[] the loaders fill, width/height/maxBounce globals of main.c:10-12)
 * and the same output (Color image[width*height], RGB8, row 0 on top,
 * main.c:100,246), ready for the unchanged image writer call at main.c:305.
 *
 * Everything is plain C: pointers and sizes, no C++ or torch types.  All
 * compute runs in hand-written CUDA kernels for sm_100a; there is NO CPU
 * fallback -- without a usable GPU every entry point returns an error code.
 */
#ifndef RTC_B200_H
#define RTC_B200_H

#include <stddef.h>
#include <stdint.h>

#include "rtc_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes (the library never calls exit(); reference exits: SURVEY 5) */
#define RTC_OK 0
#define RTC_ERR_ARG 1      /* NULL / negative / inconsistent argument           */
#define RTC_ERR_NO_GPU 2   /* no CUDA device, or not enough devices for ngpus   */
#define RTC_ERR_CUDA 3     /* a CUDA runtime call failed (message on stderr)    */
#define RTC_ERR_NCCL 4     /* libnccl missing or a NCCL call failed             */
#define RTC_ERR_NOMEM 5

/* ---- how the image is split over GPUs ------------------------- */
#define RTC_SPLIT_TILES 0   /* interleaved row bands per GPU, ncclAllGather     */
#define RTC_SPLIT_SAMPLES 1 /* disjoint sample sub-streams per GPU, float reduce */

/* ---- RenderParams.flags --------------------------------------------------- */
#define RTC_FLAG_COUNT_EXITS 1u /* fill RenderStats.tri_exit (slower kernel)     */
#define RTC_FLAG_EXACT_ONLY 2u  /* disable every conservative pre-filter: every
                                   ray-triangle pair runs the reference's own
                                   operation sequence (validation mode)          */
#define RTC_FLAG_PER_PAIR_FILTER 8u /* primary pre-filter evaluated per (pixel, triangle) pair instead of
                                      per (8x4 pixel block, triangle) + expansion of surviving blocks */
#define RTC_FLAG_TIME_KERNELS 16u /* record CUDA events around the primary and the path kernel (rtc_scene_kernel_times) */
#define RTC_FLAG_VERIFY_QUICK 32u /* path kernel: also run the exact test on EVERY (ray, triangle) pair and count in
                                    device stats word 11 the accepted pairs its conservative quick test had dropped
                                    (must be 0); needs a stats buffer; slow, for tests */
#define RTC_FLAG_REF_QUICK 64u  /* path kernel: first phase of the sweep in the reference's own arithmetic (the
                                   fallback for scenes whose stored normals do not follow the winding) */
#define RTC_FLAG_PATHS_ONLY 4u  /* no primary pre-filter: every pixel goes through the
                                   path kernel (exact arithmetic, shared-memory tiles) */
#define RTC_FLAG_SERIAL_PATHS 128u /* path kernel, first form: one lane owns one pixel and runs its samples in
                                      sequence (validation mode for the pool kernel, which evaluates a pixel's samples
                                      out of order; also taken automatically when maxBounce > 255) */

/* counters a caller can ask for; replaces the reference's printf progress */
typedef struct RenderStats {
    uint64_t rays;         /* closest-hit sweeps (calls of calculateRayCollision, raytracing.c:270) */
    uint64_t tri_pairs;    /* ray-triangle pairs tested = rays * ntris          */
    uint64_t sphere_pairs; /* ray-sphere pairs tested                            */
    uint64_t tri_exit[6];  /* pairs by reference exit: cull, det, u, v, dst, hit (RTC_FLAG_COUNT_EXITS) */
    uint64_t hit_pixels;   /* pixels whose primary ray hit (only if out_primary_id != NULL) */
    float upload_ms;       /* flatten + host->device of the scene               */
    float kernel_ms;       /* device time of the render kernels, max over GPUs  */
    float gather_ms;       /* NCCL gather / reduce + un-interleave              */
    float download_ms;     /* device->host of the image                         */
    int32_t kernels_launched;
    int32_t ngpus_used;
    uint64_t check_failures; /* device self-checks, must be 0: samples whose primary re-sweep disagreed with the
                                classification + (RTC_FLAG_VERIFY_QUICK) accepted pairs the quick test had dropped */
    /* scheduler counters of the path kernel (pool form): samples evaluated (those consumed by a pixel + speculative
     * ones that were dropped), shade passes, lanes they filled (of 32), sweep passes, rays they carried (of 64),
     * scheduler iterations.  rays/tri_pairs above count consumed samples only. */
    uint64_t pool_samples_evaluated, pool_shade_passes, pool_shade_lanes, pool_sweep_passes, pool_sweep_rays, pool_iterations;
} RenderStats;

/*
 * What the reference keeps in globals and file-statics:
 *   tris/ntris          raytracing.c's private triangles[]/triangleCount (scene.h:21-22)
 *   spheres/nspheres    scene.h:17-20 (one hard-coded sphere); used when trianglesOnly == 0
 *   width,height        main.c:10-11        maxBounce  main.c:12
 *   spp                 scene.h:26 accumulationCount (compile-time 4000 there)
 *   trianglesOnly       main.c:113,241
 */
typedef struct RenderParams {
    const Triangle *tris;
    int32_t ntris;
    const Sphere *spheres;
    int32_t nspheres;
    int32_t width, height;
    int32_t spp;
    int32_t maxBounce;
    int32_t trianglesOnly;
    int32_t ngpus;           /* <= 1: one GPU (device 0 unless RTC_DEVICE is set) */
    int32_t split_mode;      /* RTC_SPLIT_TILES | RTC_SPLIT_SAMPLES               */
    int32_t nsplit;          /* <= 1: the reference's one RNG stream per pixel.
                                G > 1: G sub-streams per pixel, sub-stream g seeded
                                x + y*W + g*W*H renders samples [g*spp/G,(g+1)*spp/G),
                                partial sums added in g order: this
                                library's own definition, NOT the reference's image.
                                RTC_SPLIT_SAMPLES needs it set (a multiple of ngpus). */
    int32_t *out_primary_id; /* optional, width*height: -1 miss, -2-k sphere k, else triangle index */
    RenderStats *stats;      /* optional */
    uint32_t flags;
} RenderParams;

/*
 * Render the whole frame into out_rgb (width*height*3 bytes, caller-owned host
 * memory, RGB, row 0 = top: the reference's Color image[]).  Blocking; uploads
 * the scene, renders on params->ngpus GPUs, gathers, downloads.  Replaces
 * reference main.c:285-302.  Returns RTC_OK or an RTC_ERR_* code.
 */
int render_cuda(const Scene *scene, const Camera *cam, uint8_t *out_rgb, const RenderParams *params);

/* free cached device buffers / NCCL communicators held between calls */
void render_cuda_shutdown(void);

/* camera basis exactly as reference main.c:252-255 (IEEE float, no FMA) */
void rtc_camera_look_at(const float origin[3], const float lookingAt[3], float fov, Camera *out);
/* normalized() of reference moremath.c:12-18, e.g. for the sun (main.c:247) */
void rtc_normalize(const float in[3], float out[3]);

/* Host code: the synthetic random-triangle scene of BASELINE.json's config 5 (splitmix64, definition in
 * rtc_api.cu / DESIGN.md; the reference has no generator).  The config uses seed 0xB200, ntris 1048576. */
void rtc_synth_scene(uint64_t seed, int32_t ntris, Triangle *out);

/* Host code: FNV-1a-64 of a byte buffer (offset 1469598103934665603, prime 1099511628211) -- the image hash of the
 * known-answer vectors (tests/golden/kat.json), so that callers can check a frame without other tooling. */
uint64_t rtc_fnv1a64(const void *data, size_t nbytes);

/* number of CUDA devices visible (0 if none / no driver) */
int rtc_device_count(void);
const char *rtc_last_error(void);
const char *rtc_version(void);

/* ---- device-resident API: scene stays in HBM, buffers are device pointers --
 * Used by multi-process drivers (one process per GPU, e.g. under torchrun):
 * each rank uploads the scene once, renders its share of the image into its
 * own device buffer on its own stream, and the ranks exchange the parts with
 * their collective library.                                                   */

typedef struct RtcScene RtcScene; /* flattened SoA triangle/sphere/material buffers on one device */

/* Flatten Triangle[] (AoS, 68 B) into 16-byte aligned SoA planes in HBM on
 * `device` ({A,n.x} {B-A,n.y} {C-A,n.z} + material table).  Replaces nothing
 * in the reference (it keeps AoS); consumes what raytracing.c:100-147 builds. */
int rtc_scene_create(int device, const Triangle *tris, int32_t ntris, const Sphere *spheres, int32_t nspheres,
                     RtcScene **out);
void rtc_scene_destroy(RtcScene *scene);
/* Re-upload (host -> device copy + flatten kernels, queued on `stream`) into an existing handle; buffers are
 * reused when large enough.  The per-frame "scene changed" path of a resident renderer. */
int rtc_scene_update(RtcScene *scene, const Triangle *tris, int32_t ntris, const Sphere *spheres, int32_t nspheres,
                     void *stream);

/* Which pixels one call renders, and where they land.  The output buffers are
 * COMPACT: out_w x out_h pixels with
 *     x = col_start + i * col_step                         (0 <= i < out_w)
 *     y = row_start + (j / band_h) * row_step + j % band_h (0 <= j < out_h)
 * so {0,1,0,1,1} over the full size is the whole image, {g*8, G*8, 0, 1, 8} is
 * rank g's share of an 8-row band interleave over G ranks, and {0,60,0,60,1}
 * is a sparse lattice.  Pixel results do not depend on the subset: every
 * pixel owns its RNG seed x + y*width (reference main.c:95).                  */
typedef struct RtcSubset {
    int32_t row_start, row_step;
    int32_t col_start, col_step;
    int32_t band_h;
} RtcSubset;

typedef struct RtcDeviceJob {
    int32_t width, height; /* of the FULL image (ray generation and seeds use these) */
    int32_t spp, maxBounce, trianglesOnly;
    int32_t nsplit;        /* total sub-streams per pixel (<= 1: reference)           */
    int32_t split_first;   /* this call renders sub-streams [split_first,             */
    int32_t split_count;   /*   split_first + split_count); 0 means all of them       */
    RtcSubset subset;
    int32_t out_w, out_h;  /* compact output size, see rtc_subset_size()              */
    uint32_t flags;
    int32_t rgb_full;      /* 1: d_rgb is the FULL width x height image (possibly another device's memory, see
                              rtc_ipc_open): every pixel is stored at its final row and column instead of its
                              compact position, so the ranks of a tile split assemble the frame by writing it --
                              no gather, no un-interleave (the reference's row threads fill `Color image[]` the
                              same way, main.c:84,100)                                 */
} RtcDeviceJob;

/* compact size of a subset of a width x height image */
void rtc_subset_size(int32_t width, int32_t height, const RtcSubset *s, int32_t *out_w, int32_t *out_h);

/*
 * Launch the render on `stream` (a cudaStream_t, NULL = default stream) of the
 * scene's device; returns after the launches are queued.  Device pointers, any
 * may be NULL:
 *   d_rgb     out_w*out_h*3 bytes   quantised image; written only when nsplit <= 1
 *   d_accum   planes*out_w*out_h*3 floats: one un-quantised partial sum per sub-stream of this call
 *             (planes = split_count, or nsplit - split_first when split_count == 0; 1 for the reference stream).
 *             Required when nsplit > 1: gather the planes of all sub-streams in g order and finish with
 *             rtc_sum_quantize_device, which adds them in that order (bit-exact for any GPU count).
 *   d_ids     out_w*out_h int32     primary-hit ids
 *   d_stats   RTC_DEVICE_STATS_WORDS uint64 counters, ADDED to (zero them first):
 *             [0] rays [1] ray-triangle pairs [2] ray-sphere pairs [3..8] pairs by reference exit
 *             (RTC_FLAG_COUNT_EXITS) [9] primary-hit pixels (d_ids != NULL) [10] pairs sent to the exact path
 *             [11] self-check failures (must stay 0) [12..17] path-kernel scheduler counters, in the order of
 *             RenderStats.pool_*
 */
#define RTC_DEVICE_STATS_WORDS 24
int rtc_render_device(const RtcScene *scene, const Scene *sky, const Camera *cam, const RtcDeviceJob *job,
                      uint8_t *d_rgb, float *d_accum, int32_t *d_ids, uint64_t *d_stats, void *stream);

/* total = plane_0 + plane_1 + ... (float adds in that order), then quantise with the reference's floatToUint
 * (moremath.c:25-30).  d_planes: nplanes*n floats; d_rgb: n bytes or NULL; d_sum: n floats or NULL. */
int rtc_sum_quantize_device(int device, const float *d_planes, int nplanes, size_t n, uint8_t *d_rgb, float *d_sum,
                            void *stream);

/* d_rgb[i] = quantise(d_accum[i]); n = number of floats */
int rtc_quantize_device(int device, const float *d_accum, uint8_t *d_rgb, size_t n, void *stream);

/* Device time of the two render kernels of the LAST rtc_render_device call on this scene that carried
 * RTC_FLAG_TIME_KERNELS: primary kernel (table preparation + classification + every primary sweep) and path kernel
 * (bounce rays, shading).  Waits for that render to finish. */
int rtc_scene_kernel_times(const RtcScene *scene, float *primary_ms, float *paths_ms);

/* number of kernels rtc_render_device launches for this job in the default mode (an estimate by formula; the count
 * of launches that actually happened is rtc_launches_total) */
int rtc_render_launch_count(const RtcScene *scene, const RtcDeviceJob *job, int want_ids);
/* kernels this library has launched since it was loaded (every launch site counts): difference two readings */
uint64_t rtc_launches_total(void);

/* A device buffer of ANOTHER process as a pointer of this one (CUDA IPC over NVLink / PCIe peer access): the owner
 * exports a 64-byte handle, peers open it and pass the pointer as d_rgb with RtcDeviceJob.rgb_full = 1.  With one
 * process per GPU this is how the ranks of a tile split write their pixels straight into the destination frame. */
#define RTC_IPC_HANDLE_BYTES 64
int rtc_ipc_export(int device, const void *d_ptr, unsigned char handle[RTC_IPC_HANDLE_BYTES]);
int rtc_ipc_open(int device, const unsigned char handle[RTC_IPC_HANDLE_BYTES], void **d_ptr);
int rtc_ipc_close(int device, void *d_ptr);
/* the same inside one process: let `device` access `peer`'s memory (idempotent) */
int rtc_enable_peer(int device, int peer);

/* parts [nparts][part_elems] -> image rows: part g holds the band_h-row bands g, g+nparts, ... of an image
 * with `height` rows of `row_elems` elements (elem_size 1 or 4 bytes).  The un-interleave step after an
 * all-gather of a tile split. */
int rtc_uninterleave_device(int device, const void *d_parts, void *d_img, int elem_size, int row_elems, int height,
                            int nparts, int band_h, size_t part_elems, void *stream);

/* One part of a band split (the rows rtc_render_device wrote for RtcSubset{part * band_h, nparts * band_h, 0, 1, band_h},
 * compact, band after band) straight into a HOST image of `height` rows of `row_bytes` bytes, asynchronously on
 * `stream`: one strided device->host copy.  With one process or thread per GPU every device delivers its own rows
 * over its own PCIe link -- the all-gather + rtc_uninterleave_device + full-frame copy are only needed when the frame
 * is wanted in DEVICE memory.  (The reference's row threads fill `Color image[]` the same way: main.c:84,100.) */
int rtc_download_bands(int device, const void *d_part, void *h_img, size_t row_bytes, int height, int nparts, int band_h,
                       int part, void *stream);

/*
 * NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * The path kernel (rtc::render_paths_kernel: every ray of every sample of every pixel against every triangle, the
 * brute-force form that RTC_FLAG_PATHS_ONLY selects in rtc_render_device) with ALL of its scratch -- the work cursor
 * and the per-scene quick records of the sweep's conservative first phase -- in ONE file-scope
 * `static __device__ __align__(256)` array of the library instead of buffers of the scene handle.  The call
 * allocates nothing: no cudaMalloc, no cudaFree, no allocating library routine.  Inputs and outputs are the
 * caller's: an existing scene handle (its flattened triangle buffers were allocated by rtc_scene_create, outside
 * the no-allocation region), sky, camera, job, and the device buffers d_rgb (required), d_ids, d_stats (optional),
 * with the meaning they have in rtc_render_device.  The image is bit-identical to rtc_render_device's.
 *
 * Limits: at most RTC_NOALLOC_MAX_TRIS triangles in the scene (more -> RTC_ERR_ARG, nothing is launched); the
 * reference's one RNG stream per pixel only (job->nsplit <= 1); job->maxBounce >= 1.  The array is
 * RTC_NOALLOC_SCRATCH_BYTES (about 144 KB) and every GPU that loads the library holds it until the process exits.
 * EVERY call on a device shares that one array and finds the previous call's contents in it: each call clears the
 * cursor words and rebuilds the quick records for its scene on `stream` before its kernel -- so calls are correct
 * back to back on one stream (or otherwise ordered on the GPU), and two calls that OVERLAP on the same GPU
 * (different streams, no dependency) RACE on the array.
 */
#define RTC_NOALLOC_MAX_TRIS 2048
#define RTC_NOALLOC_SCRATCH_BYTES 147856 /* 256 + 144 per pair of triangles, RTC_NOALLOC_MAX_TRIS / 2 + 1 pairs (checked in rtc_api.cu) */
int rtc_render_paths_noalloc(const RtcScene *scene, const Scene *sky, const Camera *cam, const RtcDeviceJob *job,
                             uint8_t *d_rgb, int32_t *d_ids, uint64_t *d_stats, void *stream);

/* Dependent-free FFMA probe: measured FP32 peak of `device` in TFLOP/s (2 flop per FFMA), best of `reps`.
 * The roofline denominator of the intersection kernel (MEASURED_PEAKS.json has no FP32 figure). */
int rtc_fp32_peak_probe(int device, int reps, double *tflops, double *ms_best);

/* Test probes (host buffers): the device versions of RandomDiretion (moremath.c:105-109; out = 3 floats per state,
 * out_states = the RNG state after the 6 draws) and getEnvironmentLight (raytracing.c:151-160; dirs/out = 3 floats
 * each) on n independent inputs, so that the FP64 islands can be compared with the reference at scale. */
int rtc_probe_random_direction(int device, const uint32_t *states, int32_t n, float *out, uint32_t *out_states);
int rtc_probe_environment_light(int device, const Scene *sky, const float *dirs, int32_t n, float *out);
/* Self-check of the normal deviate (RandomValueNormalDistrubtion, moremath.c:98-103): the device evaluates it by short
 * tailored approximations + a Ziv check and falls back to the reference's sequence when the check cannot decide.  For the
 * `count` generator states from `first_state` on (2^32 = every state the reference can reach) this compares the two, bit
 * for bit: out[0] = decided and different (must be 0), out[1] = declined (fell back), out[2] = states checked. */
int rtc_probe_normal_selfcheck(int device, uint32_t first_state, uint64_t count, uint64_t out[3]);

#ifdef __cplusplus
}
#endif

#endif /* RTC_B200_H */
