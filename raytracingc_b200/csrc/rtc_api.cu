// rtc_api.cu -- the C ABI of librtc_b200.so (include/rtc_b200.h).
//
// Host-side plumbing only: argument checks, device memory, launches, copies.
// All arithmetic of the render loop is in rtc_kernels.cuh.  There is no CPU
// rendering path here: without a CUDA device every entry point fails.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <thread>
#include <map>
#include <mutex>
#include <tuple>
#include <utility>
#include <vector>

#include <cuda_runtime.h>

#include "rtc_b200.h"
#include "rtc_kernels.cuh"
#include "rtc_fast.cuh"
#include "rtc_paths.cuh"
#include "rtc_pool.cuh"
#include "rtc_sky.cuh"
#include "rtc_skyblock.cuh"

static_assert(sizeof(Triangle) == 68, "Triangle must match reference raytracing.h:41-45");
static_assert(sizeof(Sphere) == 36, "Sphere must match reference raytracing.h:34-39");
static_assert(sizeof(Scene) == 56, "Scene must match reference raytracing.h:7-11");
static_assert(sizeof(Camera) == 52, "Camera is origin, ex, ey, ez, fov");

namespace {

thread_local char g_err[512] = "";

// kernels launched by this library since it was loaded (every <<< >>> below is followed by count_launch())
std::atomic<unsigned long long> g_launches{0};
inline void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    fprintf(stderr, "rtc_b200: %s\n", g_err);
    return code;
}

#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return fail(RTC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

rtc::f3 to3(const vec3 &v) { return rtc::f3{v.x, v.y, v.z}; }

// host copy of reference normalized(), moremath.c:7-18 (built with -ffp-contract=off, no FMA on x86-64 baseline)
void host_normalize(const float in[3], float out[3])
{
    float len = (float)sqrt((double)(in[0] * in[0] + in[1] * in[1] + in[2] * in[2]));
    float inv = (float)(1. / (double)len);
    out[0] = in[0] * inv;
    out[1] = in[1] * inv;
    out[2] = in[2] * inv;
}

void host_cross(const float u[3], const float v[3], float out[3])
{
    out[0] = u[1] * v[2] - u[2] * v[1];
    out[1] = u[2] * v[0] - u[0] * v[2];
    out[2] = u[0] * v[1] - u[1] * v[0];
}

} // namespace

struct RtcScene {
    int device = 0;
    int ntris = 0, nsph = 0;
    float4 *geo = nullptr, *m0 = nullptr;
    float *m1 = nullptr;
    float4 *s0 = nullptr, *s1 = nullptr;
    float *s2 = nullptr;
    float *raw = nullptr; // staging of the caller's AoS records
    size_t raw_bytes = 0;
    int cap_tris = 0, cap_sph = 0;
    float4 *tab = nullptr; // per-frame edge-function table of the primary pre-filter (4 x float4 per triangle)
    int cap_tab = 0;
    int max_smem_optin = 0;
    int nsm = 0;
    int *list = nullptr;          // per-frame work list of primary-hit pixels (sky kernel -> path kernel)
    size_t cap_list = 0;
    float2 *rec = nullptr;        // primary-hit record of each listed pixel
    int4 *btask = nullptr;        // blocks with candidate triangles (block kernel -> block-task kernel)
    float *sphq = nullptr;        // per-frame sphere bound table (8 floats per sphere)
    int cap_sphq = 0;
    cudaStream_t aux = nullptr;   // the block-task kernel runs here, beside the path kernel
    cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
    size_t cap_btask = 0;
    cudaEvent_t kev[3] = {nullptr, nullptr, nullptr}; // RTC_FLAG_TIME_KERNELS: before primary, between, after paths
    bool kev_valid = false;
    unsigned *counters = nullptr; // [0] list length, [1] path-kernel work cursor, [2] block tasks, [3] their cursor,
                                  // [4] pool kernel: scheduler watchdog tripped
    float4 *qrec = nullptr;       // quick records of the path kernel's first phase (rtc_paths.cuh)
    int q_layout = 0;             // layout qrec currently holds: 1 per triangle, 2 per pair (packed), 0 none
    int cap_q = 0;
    float q_obound = 0.f;         // origin bound the records were built for; 0 = not built for the current scene
    float bound_abs = 0.f;        // max |coordinate| over triangle vertices and sphere extents (host scan at upload)
    bool quick_ok = false;        // the stored normals follow the winding closely enough for the quick records to pay
    // pipelined device->host copy of render_cuda (one-shot, consumed by the next render): see RenderArgs::prog
    unsigned *prog = nullptr;
    int prog_rows = 0, prog_chunks = 0;
    // length of the last frame's work list, copied to a page-locked word behind the block-task kernel (off the critical
    // path).  Only a HINT for which form of the path kernel to queue first (both stay correct whatever it says).
    unsigned *h_list_len = nullptr;
    bool list_len_seen = false;
    cudaEvent_t sky_ev = nullptr; // recorded when every pixel no later kernel rewrites is final (sky pixels)
    bool sky_recorded = false;
};

extern "C" {

const char *rtc_last_error(void) { return g_err; }
const char *rtc_version(void) { return "rtc_b200 0.1 (sm_100a)"; }

int rtc_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

void rtc_normalize(const float in[3], float out[3]) { host_normalize(in, out); }

uint64_t rtc_fnv1a64(const void *data, size_t nbytes)
{
    const unsigned char *p = static_cast<const unsigned char *>(data);
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < nbytes; ++i)
        h = (h ^ p[i]) * 1099511628211ull;
    return h;
}

// reference main.c:252-255
void rtc_camera_look_at(const float origin[3], const float lookingAt[3], float fov, Camera *out)
{
    float d[3] = {lookingAt[0] - origin[0], lookingAt[1] - origin[1], lookingAt[2] - origin[2]};
    float ez[3], ex[3], ey[3], t[3];
    const float up[3] = {0, -1, 0};
    host_normalize(d, ez);
    host_cross(ez, up, t);
    host_normalize(t, ex);
    host_cross(ez, ex, t);
    host_normalize(t, ey);
    out->origin = vec3{origin[0], origin[1], origin[2]};
    out->ex = vec3{ex[0], ex[1], ex[2]};
    out->ey = vec3{ey[0], ey[1], ey[2]};
    out->ez = vec3{ez[0], ez[1], ez[2]};
    out->fov = fov;
}

// ---- synthetic scene of BASELINE.json's config 5 (the reference has no generator; SURVEY.md 8d defines it) -----
// splitmix64 stream; per triangle, in this order: centre (3 draws, uniform in [-8,8] x [-6,2] x [-2,14], in front of
// the reference's default camera), edge vectors e1, e2 (6 draws, uniform in [-0.08,0.08]^3), colour (3 draws,
// uniform in [0.2,1]), emission (1 draw: 4 for one value in 64, else 0), smoothness (1 draw, uniform in [0,1]).
// A = c, B = c + e1, C = c + e2 in float; normal = normalized(cross(B - A, C - A)) with the reference's float
// operations (the rule of raytracing.c:24).  A draw is u = (x >> 11) * 2^-53; value = (float)(lo + (hi - lo) * u).
void rtc_synth_scene(uint64_t seed, int32_t ntris, Triangle *out)
{
    uint64_t state = seed;
    auto next = [&state]() {
        uint64_t z = (state += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    };
    auto uni = [&next](double lo, double hi) { return (float)(lo + (hi - lo) * ((double)(next() >> 11) * 0x1p-53)); };
    for (int32_t i = 0; i < ntris; ++i) {
        Triangle &t = out[i];
        const float cx = uni(-8, 8), cy = uni(-6, 2), cz = uni(-2, 14);
        const float e1[3] = {uni(-0.08, 0.08), uni(-0.08, 0.08), uni(-0.08, 0.08)};
        const float e2[3] = {uni(-0.08, 0.08), uni(-0.08, 0.08), uni(-0.08, 0.08)};
        t.posA = vec3{cx, cy, cz};
        t.posB = vec3{cx + e1[0], cy + e1[1], cz + e1[2]};
        t.posC = vec3{cx + e2[0], cy + e2[1], cz + e2[2]};
        const float ab[3] = {t.posB.x - t.posA.x, t.posB.y - t.posA.y, t.posB.z - t.posA.z};
        const float ac[3] = {t.posC.x - t.posA.x, t.posC.y - t.posA.y, t.posC.z - t.posA.z};
        float n[3], nn[3];
        host_cross(ab, ac, n);
        host_normalize(n, nn);
        t.normal = vec3{nn[0], nn[1], nn[2]};
        t.mat.color = vec3{uni(0.2, 1), uni(0.2, 1), uni(0.2, 1)};
        t.mat.emissionStrength = (next() & 63u) == 0 ? 4.0f : 0.0f;
        t.mat.smoothness = uni(0, 1);
    }
}

void rtc_subset_size(int32_t width, int32_t height, const RtcSubset *s, int32_t *out_w, int32_t *out_h)
{
    int cs = s->col_step > 0 ? s->col_step : 1, rs = s->row_step > 0 ? s->row_step : 1;
    int bh = s->band_h > 0 ? s->band_h : 1;
    int w = s->col_start < width ? (width - s->col_start + cs - 1) / cs : 0;
    int h = 0;
    for (int y0 = s->row_start; y0 < height; y0 += rs) {
        int rows = height - y0 < bh ? height - y0 : bh;
        h += rows;
        if (rs < bh) { // overlapping bands are not a valid subset; treat as contiguous
            h = height - s->row_start;
            break;
        }
    }
    *out_w = w;
    *out_h = h;
}

void rtc_scene_destroy(RtcScene *sc)
{
    if (!sc)
        return;
    cudaSetDevice(sc->device);
    cudaFree(sc->geo);
    cudaFree(sc->m0);
    cudaFree(sc->m1);
    cudaFree(sc->s0);
    cudaFree(sc->s1);
    cudaFree(sc->s2);
    cudaFree(sc->raw);
    cudaFree(sc->tab);
    cudaFree(sc->list);
    cudaFree(sc->rec);
    cudaFree(sc->btask);
    cudaFree(sc->sphq);
    cudaFree(sc->qrec);
    cudaFree(sc->counters);
    cudaFreeHost(sc->h_list_len);
    for (auto &e : sc->kev)
        if (e)
            cudaEventDestroy(e);
    if (sc->sky_ev)
        cudaEventDestroy(sc->sky_ev);
    if (sc->fork_ev)
        cudaEventDestroy(sc->fork_ev);
    if (sc->join_ev)
        cudaEventDestroy(sc->join_ev);
    if (sc->aux)
        cudaStreamDestroy(sc->aux);
    delete sc;
}

static int scene_reserve(RtcScene *sc, int ntris, int nsph)
{
    CU(cudaSetDevice(sc->device));
    if (ntris > sc->cap_tris || !sc->geo) {
        size_t nt = ntris > 0 ? ntris : 1;
        cudaFree(sc->geo); cudaFree(sc->m0); cudaFree(sc->m1);
        sc->geo = sc->m0 = nullptr;
        sc->m1 = nullptr;
        CU(cudaMalloc(&sc->geo, sizeof(float4) * 3 * nt));
        CU(cudaMalloc(&sc->m0, sizeof(float4) * nt));
        CU(cudaMalloc(&sc->m1, sizeof(float) * nt));
        sc->cap_tris = (int)nt;
    }
    if (nsph > sc->cap_sph || !sc->s0) {
        size_t ns = nsph > 0 ? nsph : 1;
        cudaFree(sc->s0); cudaFree(sc->s1); cudaFree(sc->s2);
        sc->s0 = sc->s1 = nullptr;
        sc->s2 = nullptr;
        CU(cudaMalloc(&sc->s0, sizeof(float4) * ns));
        CU(cudaMalloc(&sc->s1, sizeof(float4) * ns));
        CU(cudaMalloc(&sc->s2, sizeof(float) * ns));
        sc->cap_sph = (int)ns;
    }
    size_t need = sizeof(Triangle) * (size_t)(ntris > 0 ? ntris : 1) + sizeof(Sphere) * (size_t)(nsph > 0 ? nsph : 1);
    if (need > sc->raw_bytes) {
        cudaFree(sc->raw);
        sc->raw = nullptr;
        CU(cudaMalloc(&sc->raw, need));
        sc->raw_bytes = need;
    }
    return RTC_OK;
}

// Host scan at upload: the coordinate bound the quick records (rtc_paths.cuh) are built for, and whether they pay:
// a record culls by the sign of d.(AC x AB), which needs the stored normal to be (nearly) parallel to the
// geometric one; a triangle whose normal is off by more than ~0.25 degree is merely kept more often (never wrong),
// and a scene with more than 1/16 of such triangles uses the reference-arithmetic quick test instead.
static void scan_scene(RtcScene *sc, const Triangle *tris, int ntris, const Sphere *spheres, int nsph)
{
    float bound = 0.f;
    long long off = 0;
    for (int i = 0; i < ntris; ++i) {
        const float *t = reinterpret_cast<const float *>(&tris[i]);
        for (int k = 0; k < 9; ++k)
            bound = std::fmax(bound, std::fabs(t[k]));
        const double ab[3] = {(double)t[3] - t[0], (double)t[4] - t[1], (double)t[5] - t[2]};
        const double ac[3] = {(double)t[6] - t[0], (double)t[7] - t[1], (double)t[8] - t[2]};
        const double N[3] = {ac[1] * ab[2] - ac[2] * ab[1], ac[2] * ab[0] - ac[0] * ab[2], ac[0] * ab[1] - ac[1] * ab[0]};
        const double nn = t[9] * N[0] + t[10] * N[1] + t[11] * N[2], NN = N[0] * N[0] + N[1] * N[1] + N[2] * N[2];
        const double n2 = (double)t[9] * t[9] + (double)t[10] * t[10] + (double)t[11] * t[11];
        // sin^2 of the angle between n and N: 1 - (n.N)^2 / (|n|^2 |N|^2)
        if (!(NN > 0.0) || !(n2 > 0.0) || !(1.0 - nn * nn / (NN * n2) < 2e-5))
            ++off;
    }
    for (int i = 0; i < nsph; ++i) {
        const float *t = reinterpret_cast<const float *>(&spheres[i]);
        for (int k = 0; k < 3; ++k)
            bound = std::fmax(bound, std::fabs(t[k]) + std::fabs(t[3]));
    }
    sc->bound_abs = bound;
    sc->quick_ok = off * 16 <= (long long)ntris;
    sc->q_obound = 0.f;
}

// copy the caller's AoS arrays to the device and flatten them there (one-shot kernels), on `st`
static int scene_upload(RtcScene *sc, const Triangle *tris, int ntris, const Sphere *spheres, int nsph,
                        cudaStream_t st)
{
    int rc = scene_reserve(sc, ntris, nsph);
    if (rc)
        return rc;
    sc->ntris = ntris;
    sc->nsph = nsph;
    scan_scene(sc, tris, ntris, spheres, nsph);
    float *raw_t = sc->raw;
    float *raw_s = reinterpret_cast<float *>(reinterpret_cast<char *>(sc->raw) +
                                             sizeof(Triangle) * (size_t)(ntris > 0 ? ntris : 1));
    if (nsph > 0) {
        CU(cudaMemcpyAsync(raw_s, spheres, sizeof(Sphere) * (size_t)nsph, cudaMemcpyHostToDevice, st));
        rtc::flatten_spheres_kernel<<<(nsph + 127) / 128, 128, 0, st>>>(raw_s, nsph, sc->s0, sc->s1, sc->s2);
        count_launch();
        CU(cudaGetLastError());
    }
    if (ntris > 0) {
        CU(cudaMemcpyAsync(raw_t, tris, sizeof(Triangle) * (size_t)ntris, cudaMemcpyHostToDevice, st));
        rtc::flatten_kernel<<<(ntris + 127) / 128, 128, 0, st>>>(raw_t, ntris, sc->geo, sc->m0, sc->m1);
        count_launch();
        CU(cudaGetLastError());
    }
    return RTC_OK;
}

int rtc_scene_create(int device, const Triangle *tris, int32_t ntris, const Sphere *spheres, int32_t nspheres,
                     RtcScene **out)
{
    if (!out || ntris < 0 || nspheres < 0 || (ntris > 0 && !tris) || (nspheres > 0 && !spheres))
        return fail(RTC_ERR_ARG, "rtc_scene_create: bad arguments");
    int ndev = rtc_device_count();
    if (ndev <= 0)
        return fail(RTC_ERR_NO_GPU, "no CUDA device: librtc_b200 has no CPU fallback");
    if (device < 0 || device >= ndev)
        return fail(RTC_ERR_NO_GPU, "device %d requested, %d visible", device, ndev);
    RtcScene *sc = new RtcScene;
    sc->device = device;
    int rc = scene_upload(sc, tris, ntris, spheres, nspheres, 0);
    if (rc == RTC_OK && cudaDeviceSynchronize() != cudaSuccess)
        rc = fail(RTC_ERR_CUDA, "scene upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc) {
        rtc_scene_destroy(sc);
        return rc;
    }
    *out = sc;
    return RTC_OK;
}

} // extern "C"

extern "C" int rtc_scene_update(RtcScene *sc, const Triangle *tris, int32_t ntris, const Sphere *spheres,
                                int32_t nspheres, void *stream)
{
    if (!sc || ntris < 0 || nspheres < 0 || (ntris > 0 && !tris) || (nspheres > 0 && !spheres))
        return fail(RTC_ERR_ARG, "rtc_scene_update: bad arguments");
    return scene_upload(sc, tris, ntris, spheres, nspheres, static_cast<cudaStream_t>(stream));
}

// ---- device-resident render -------------------------------------------------------

namespace {

int check_job(const RtcScene *sc, const Scene *sky, const Camera *cam, const RtcDeviceJob *job)
{
    if (!sc || !sky || !cam || !job)
        return fail(RTC_ERR_ARG, "render: NULL argument");
    if (job->width <= 0 || job->height <= 0 || job->spp <= 0 || job->maxBounce < 0)
        return fail(RTC_ERR_ARG, "render: width/height/spp must be > 0, maxBounce >= 0 (got %d x %d, spp %d, b %d)",
                    job->width, job->height, job->spp, job->maxBounce);
    if (job->height / 2 == 0)
        return fail(RTC_ERR_ARG, "render: height must be >= 2 (the reference divides by height/2, main.c:88)");
    if (job->out_w < 0 || job->out_h < 0)
        return fail(RTC_ERR_ARG, "render: negative output size");
    if (job->subset.row_start < 0 || job->subset.col_start < 0)
        return fail(RTC_ERR_ARG, "render: negative subset origin");
    {
        // the compact size must be the subset's: a larger out_h would write outside the caller's buffers
        int32_t w = 0, h = 0;
        rtc_subset_size(job->width, job->height, &job->subset, &w, &h);
        if (w != job->out_w || h != job->out_h)
            return fail(RTC_ERR_ARG, "render: out_w x out_h = %d x %d but the subset has %d x %d pixels (rtc_subset_size)",
                        job->out_w, job->out_h, w, h);
    }
    {
        const long long G = job->nsplit > 1 ? job->nsplit : 1;
        const long long planes = job->split_count > 0 ? job->split_count : G - job->split_first;
        if (planes > 65535)
            return fail(RTC_ERR_ARG, "render: at most 65535 sub-streams per call (got %lld)", planes);
        if ((long long)job->out_w * job->out_h * (planes > 0 ? planes : 1) >= (1ll << 31))
            return fail(RTC_ERR_ARG, "render: out_w * out_h * sub-streams must stay below 2^31 (32-bit task counters)");
    }
    return RTC_OK;
}

rtc::SkyParams make_sky(const Scene *sky)
{
    rtc::SkyParams k;
    k.sun = to3(sky->normalizedSunDirection);
    k.horizon = to3(sky->skyColorHorizon);
    k.zenith = to3(sky->skyColorZenith);
    k.ground = to3(sky->groundColor);
    k.sunFocus = sky->sunFocus;
    k.sunIntensity = sky->sunIntensity;
    k.sun_skip_ok = (sky->sunFocus >= 0.0f && std::isfinite(sky->sunIntensity)) ? 1 : 0;
    return k;
}

rtc::RenderArgs make_args(const RtcScene *sc, const Scene *sky, const Camera *cam, const RtcDeviceJob *job)
{
    rtc::RenderArgs a;
    memset(&a, 0, sizeof a);
    a.sc.geo = sc->geo; a.sc.m0 = sc->m0; a.sc.m1 = sc->m1;
    a.sc.ntris = sc->ntris;
    a.sc.s0 = sc->s0; a.sc.s1 = sc->s1; a.sc.s2 = sc->s2;
    a.sc.nsph = sc->nsph;
    a.sky = make_sky(sky);
    a.origin = to3(cam->origin);
    a.ex = to3(cam->ex);
    a.ey = to3(cam->ey);
    a.ez = to3(cam->ez);
    a.fov = cam->fov;
    a.W = job->width;
    a.H = job->height;
    a.spp = job->spp;
    a.maxBounce = job->maxBounce;
    a.trianglesOnly = job->trianglesOnly;
    a.nsplit = job->nsplit > 1 ? job->nsplit : 1;
    a.split_first = job->split_first;
    a.row_start = job->subset.row_start;
    a.row_step = job->subset.row_step > 0 ? job->subset.row_step : 1;
    a.col_start = job->subset.col_start;
    a.col_step = job->subset.col_step > 0 ? job->subset.col_step : 1;
    a.band_h = job->subset.band_h > 0 ? job->subset.band_h : 1;
    a.out_w = job->out_w;
    a.out_h = job->out_h;
    a.rgb_full = job->rgb_full ? 1 : 0;
    return a;
}

// ---- fast-path eligibility and per-frame constants -----------------------------------

double dot3d(const vec3 &a, const vec3 &b) { return (double)a.x * b.x + (double)a.y * b.y + (double)a.z * b.z; }

// The pre-filter's error analysis (rtc_fast.cuh) assumes the camera basis main.c:252-255 builds:
// orthonormal up to rounding.  Any other basis is rendered by the exact kernel.
bool camera_is_orthonormal(const Camera *c)
{
    const double tol = 1e-3;
    const vec3 *e[3] = {&c->ex, &c->ey, &c->ez};
    for (int i = 0; i < 3; ++i) {
        if (!(fabs(dot3d(*e[i], *e[i]) - 1.0) < tol))
            return false;
        for (int j = i + 1; j < 3; ++j)
            if (!(fabs(dot3d(*e[i], *e[j])) < tol))
                return false;
    }
    return std::isfinite(c->fov) && std::isfinite(c->origin.x) && std::isfinite(c->origin.y) &&
           std::isfinite(c->origin.z);
}

// env override helpers (test / tuning knobs; unset in production)
// cudaFuncSetAttribute + occupancy query of a kernel at a dynamic shared-memory size.  Both runtime calls cost
// several microseconds and a frame makes a dozen of them, so the opt-in limit is only ever raised (it is sticky per
// function) and the occupancy is remembered per (device, kernel, size).
int env_int(const char *name, int dflt);

int kernel_occupancy(int device, const void *kern, int threads, size_t smem, int *per_sm)
{
    using FnKey = std::pair<int, const void *>;
    static std::map<FnKey, size_t> limit;                      // opt-in shared memory currently set
    static std::map<std::pair<FnKey, size_t>, int> occupancy;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    const FnKey fk{device, kern};
    auto li = limit.find(fk);
    if (li == limit.end() || li->second < smem) {
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (env_int("RTC_CARVEOUT", 0) != 0) // A/B knob: every render kernel asks for the largest shared-memory carve-out, so that
                                             // kernels of different footprints can follow each other without reconfiguring the SMs
            CU(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
        limit[fk] = smem;
    }
    auto oi = occupancy.find({fk, smem});
    if (oi == occupancy.end()) {
        int n = 0;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, threads, smem));
        oi = occupancy.emplace(std::make_pair(fk, smem), n).first;
    }
    *per_sm = oi->second;
    return RTC_OK;
}

int env_int(const char *name, int dflt)
{
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}

// pixels per lane of the sky kernel, 0 = the filtered path is not applicable (exact kernels only)
int fast_path_R(const RtcScene *sc, const Camera *cam, const RtcDeviceJob *job)
{
    (void)sc;
    if (job->flags & (RTC_FLAG_EXACT_ONLY | RTC_FLAG_COUNT_EXITS | RTC_FLAG_PATHS_ONLY))
        return 0;
    if (job->maxBounce < 1 || !camera_is_orthonormal(cam))
        return 0;
    if ((long long)job->width * job->height >= (1ll << 31))
        return 0;
    int G = job->nsplit > 1 ? job->nsplit : 1;
    if (job->spp < G)
        return 0; // some sub-stream has no samples: no classification sample to hang the work list on
    const char *e = getenv("RTC_FAST_R");
    int R = e ? atoi(e) : -1;
    if (e && R == 0)
        return 0;
    if (R != 1 && R != 2 && R != 4 && R != 8 && R != 16) {
        long long px = (long long)job->out_w * job->out_h;
        R = px >= 4000000 ? 8 : (px >= 1000000 ? 4 : 2);
    }
    return R;
}

int ensure_scratch(RtcScene *sc, size_t npix, size_t nbtask = 0)
{
    if (nbtask > sc->cap_btask) {
        CU(cudaFree(sc->btask));
        sc->btask = nullptr;
        CU(cudaMalloc(&sc->btask, sizeof(int4) * nbtask));
        sc->cap_btask = nbtask;
    }
    if (!sc->counters)
        CU(cudaMalloc(&sc->counters, 8 * sizeof(unsigned)));
    if (!sc->h_list_len) {
        CU(cudaMallocHost(&sc->h_list_len, sizeof(unsigned)));
        *sc->h_list_len = 0u;
    }
    if (npix > sc->cap_list || !sc->list) {
        CU(cudaFree(sc->list));
        sc->list = nullptr;
        CU(cudaMalloc(&sc->list, sizeof(int) * (npix ? npix : 1)));
        CU(cudaFree(sc->rec));
        sc->rec = nullptr;
        CU(cudaMalloc(&sc->rec, sizeof(float2) * (npix ? npix : 1)));
        sc->cap_list = npix;
    }
    if (!sc->nsm) {
        CU(cudaDeviceGetAttribute(&sc->nsm, cudaDevAttrMultiProcessorCount, sc->device));
        CU(cudaDeviceGetAttribute(&sc->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, sc->device));
    }
    return RTC_OK;
}

// quick records of the path kernels' first phase (rtc_paths.cuh), built once per scene and again only when a ray
// origin leaves the bound they were built for or the layout changes
int ensure_quick_records(RtcScene *sc, const rtc::RenderArgs &a, int mode, bool has_rec, cudaStream_t st)
{
    // bounce rays start on the scene's surfaces; only when the path kernel sweeps the primary rays too does the camera count
    float need = sc->bound_abs;
    if (!has_rec)
        need = std::fmax(need, std::fmax(std::fabs(a.origin.x), std::fmax(std::fabs(a.origin.y), std::fabs(a.origin.z))));
    need = need * 1.001f + 1e-6f;
    if (sc->ntris > sc->cap_q || !sc->qrec) {
        cudaFree(sc->qrec);
        sc->qrec = nullptr;
        // room for either layout: 4 float4 per triangle, or QPAIR_F4 per pair of triangles
        CU(cudaMalloc(&sc->qrec, sizeof(float4) * (rtc::path_q_f4(5, sc->ntris) + rtc::QPAIR_F4)));
        sc->cap_q = sc->ntris > 0 ? sc->ntris : 1;
        sc->q_obound = 0.f;
    }
    const int q_layout = rtc::mode_layout(mode);
    if ((!(need <= sc->q_obound) || sc->q_layout != q_layout) && sc->ntris > 0) {
        const float ob = sc->q_layout != q_layout && need <= sc->q_obound ? sc->q_obound : need * 1.25f;
        rtc::quick_records_kernel<<<(sc->ntris + 127) / 128, 128, 0, st>>>(sc->geo, sc->ntris, ob, sc->qrec, q_layout);
        count_launch();
        CU(cudaGetLastError());
        sc->q_obound = ob;
        sc->q_layout = q_layout;
    }
    return RTC_OK;
}

// The path kernel, pool form (rtc_pool.cuh): the default for every frame it supports.
int launch_pool(RtcScene *sc, const rtc::RenderArgs &a, const int *list, const float2 *rec, int planes, int mode,
                unsigned max_tasks, cudaStream_t st)
{
    rtc::PoolWork w;
    memset(&w, 0, sizeof w);
    if (mode != 0) {
        int rc_q = ensure_quick_records(sc, a, mode, rec != nullptr, st);
        if (rc_q)
            return rc_q;
        w.qrec = sc->qrec;
        w.obound = sc->q_obound;
    }
    w.list = list;
    w.rec = rec;
    w.list_count = sc->counters;
    w.cursor = sc->counters + 1;
    w.error = sc->counters + 4;
    w.planes = planes;
    w.max_tasks = max_tasks;
    w.spec = env_int("RTC_NO_SPEC", 0) == 0;
    w.ctx_share_div = std::max(1, env_int("RTC_POOL_SHARE_DIV", 2));
    w.tile = env_int("RTC_PATH_TILE", rtc::PATH_TILE);
    if (w.tile < 1 || w.tile > rtc::PATH_TILE)
        w.tile = rtc::PATH_TILE;
    if (rtc::mode_packed(mode) && (w.tile & 1))
        w.tile += 1; // pair records: a tile starts on an even triangle (PATH_TILE is even)
    // configuration by the size of the resident scene: two 8-warp CTAs per SM, one 16-warp CTA, or streamed tiles
    const size_t budget = (size_t)sc->max_smem_optin; // per CTA; an SM holds this + 1 KB per CTA as well
    const size_t res8 = rtc::pool_smem_bytes(sc->ntris, false, mode, rtc::POOL_WARPS_SMALL);
    const size_t res16 = rtc::pool_smem_bytes(sc->ntris, false, mode, rtc::POOL_WARPS_LARGE);
    const size_t sm_total = budget + 1024; // an SM's shared memory; every resident CTA also reserves 1 KB of it
    int cfg = (res8 + 1024) * rtc::POOL_MINB_SMALL <= sm_total ? 0 : (res16 <= budget ? 1 : 2);
    const int force = env_int("RTC_POOL_CFG", -1);
    if (env_int("RTC_FORCE_STREAM", 0) != 0)
        cfg = 2;
    else if (force >= cfg && force <= 2)
        cfg = force; // a larger configuration than needed (tests, A/B)
    const int warps = cfg == 0 ? rtc::POOL_WARPS_SMALL : rtc::POOL_WARPS_LARGE;
    const size_t smem = rtc::pool_smem_bytes(sc->ntris, cfg == 2, mode, warps);
    using PoolKernel = void (*)(const rtc::RenderArgs, const rtc::PoolWork);
    // [cfg][mode slot]: mode 0 (reference arithmetic), 3 (packed quick records), 4 (3 + verification)
#define RTC_POOL_ROW(S_, W_) \
    {rtc::render_pool_kernel<S_, 0, W_>, rtc::render_pool_kernel<S_, 3, W_>, rtc::render_pool_kernel<S_, 4, W_>}
    static const PoolKernel kernels[3][3] = {RTC_POOL_ROW(false, rtc::POOL_WARPS_SMALL), RTC_POOL_ROW(false, rtc::POOL_WARPS_LARGE),
                                             RTC_POOL_ROW(true, rtc::POOL_WARPS_LARGE)};
#undef RTC_POOL_ROW
    const PoolKernel kern = kernels[cfg][mode == 0 ? 0 : (mode == 3 ? 1 : 2)];
    int per_sm = 0;
    int rc_o = kernel_occupancy(sc->device, reinterpret_cast<const void *>(kern), 32 * warps, smem, &per_sm);
    if (rc_o)
        return rc_o;
    if (per_sm < 1)
        return fail(RTC_ERR_CUDA, "pool kernel does not fit on an SM (smem %zu)", smem);
    const long long tasks = (long long)a.out_w * a.out_h * planes; // upper bound when a list is used
    const long long want = (tasks + warps - 1) / warps;           // at least one pixel per warp
    const int blocks = (int)std::min<long long>((long long)sc->nsm * per_sm, want > 0 ? want : 1);
    // a scheduler iteration runs a pass or hands out work; far more of them than a warp could ever need means a bug:
    // the kernel then stops and reports instead of hanging the device
    const double bound = (double)tasks * (double)a.spp * (2.0 * a.maxBounce + 4.0) + 1e4;
    w.watchdog = (unsigned long long)std::min(bound, 4e7);
    kern<<<blocks, 32 * warps, smem, st>>>(a, w);
    count_launch();
    CU(cudaGetLastError());
    return RTC_OK;
}

// the path kernel: persistent grid over a work list (list == nullptr: every pixel of the subset)
int launch_paths(RtcScene *sc, const rtc::RenderArgs &a, const int *list, const float2 *rec, int planes, int write_ids,
                 unsigned flags, cudaStream_t st)
{
    rtc::PathWork w;
    // phase 1 of the sweep: quick records (1), + verification against the exact test (2), or the reference's arithmetic (0)
    const bool ref_quick = (flags & RTC_FLAG_REF_QUICK) != 0 || env_int("RTC_REF_QUICK", 0) != 0;
    // quick records two triangles at a time with packed FP32 (modes 3, 4) unless RTC_QUICK_PACKED=0 (modes 1, 2)
    // RTC_QUICK_FORM: 1 one triangle at a time (modes 1, 2), 3 two triangles at a time with packed FP32 (3, 4; default).
    // (RTC_QUICK_PACKED=0 is another spelling of form 1.)
    int form = env_int("RTC_QUICK_FORM", env_int("RTC_QUICK_PACKED", 1) != 0 ? 3 : 1);
    if (form != 1 && form != 3)
        form = 3;
    const int mode = ref_quick ? 0 : ((flags & RTC_FLAG_VERIFY_QUICK) ? form + 1 : (sc->quick_ok ? form : 0));
    if (rtc::mode_verify(mode) && !a.stats)
        return fail(RTC_ERR_ARG, "RTC_FLAG_VERIFY_QUICK needs a stats buffer");
    // Two forms of the path kernel.  The pool form (samples of a pixel out of order, full warps in every pass, two
    // rays per lane) is the one for SHORT work lists -- a frame split over several GPUs, a small image, many samples on
    // few pixels -- where the serial form runs out of pixels to keep its lanes busy; with more listed pixels than the
    // serial form has lanes, its register-resident state and absence of scheduling code win by a few percent (C4, path
    // kernel, one B200: 3.43 ms against 3.6 ms; on a 1/8 share 1.07 ms against 0.66 ms).  The list's length is only
    // known on the device, so both are launched and the one the list is not meant for returns at once.
    // RTC_PATHS=pool|serial (or RTC_FLAG_SERIAL_PATHS) forces one form.  The pool form needs the primary hits
    // (list + rec), packed or reference-arithmetic quick tests, and paths no longer than its 8-bit counters.
    const char *force = getenv("RTC_PATHS");
    const bool pool_ok = list != nullptr && rec != nullptr && a.maxBounce <= rtc::POOL_MAX_BOUNCE &&
                         (mode == 0 || mode == 3 || mode == 4);
    const bool want_serial = (flags & RTC_FLAG_SERIAL_PATHS) != 0 || env_int("RTC_PATHS_V1", 0) != 0 ||
                             (force && strcmp(force, "serial") == 0);
    const bool want_pool = force && strcmp(force, "pool") == 0;
    unsigned split_at = 0; // lists shorter than this go to the pool form
    bool pool_deferred = false, pool_only = false;
    // a streamed scene (tiles of a large scene through shared memory, CTA-synchronous sweeps) stays with the serial form
    // unless the pool is asked for: a sweep then costs a pass over the whole scene, and the serial form's cooperative
    // sweep of few rays (S lanes per ray) is the better answer to a sparse second bounce (measured on the 1 M-triangle
    // lattice: 2.9e11 against 1.3e11 pairs/s)
    const bool streamed = env_int("RTC_FORCE_STREAM", 0) != 0 ||
                          rtc::pool_smem_bytes(sc->ntris, false, mode, rtc::POOL_WARPS_LARGE) > (size_t)sc->max_smem_optin;
    if (pool_ok && !want_serial && (!streamed || want_pool)) {
        const unsigned serial_lanes = (unsigned)sc->nsm * 6u * rtc::PATH_THREADS;
        split_at = want_pool ? 0xffffffffu : (unsigned)((unsigned long long)serial_lanes * (unsigned)std::max(0, env_int("RTC_POOL_MAX_FACTOR", 125)) / 100u);
        // Which of the two is queued first matters (C4, one B200: an empty pool launch in front of the serial form costs
        // 0.07 ms -- its large shared-memory footprint breaks the overlap with the block-task kernel --, an empty serial
        // launch in front of the pool form 0.11 ms on a 1/8 share): the form that is EXPECTED to run goes first.  With
        // fewer pixels than the threshold the list is certainly short and the serial form is not launched at all;
        // otherwise the previous frame's list length decides (a hint, read without synchronisation; the first frame
        // queues the pool form first).  RTC_POOL_LAST=0|1 forces the order.
        const long long max_tasks = (long long)a.out_w * a.out_h * planes;
        pool_only = !want_pool && max_tasks < (long long)split_at;
        const bool hint_long = sc->list_len_seen && planes == 1 && sc->h_list_len != nullptr &&
                               *(volatile unsigned *)sc->h_list_len >= split_at;
        pool_deferred = !want_pool && !pool_only && env_int("RTC_POOL_LAST", hint_long ? 1 : 0) != 0;
        if (!pool_deferred) {
            int rc_p = launch_pool(sc, a, list, rec, planes, mode, split_at, st);
            if (rc_p || want_pool || pool_only)
                return rc_p;
        }
    }
    w.qrec = nullptr;
    w.obound = 0.f;
    if (mode != 0) {
        int rc_q = ensure_quick_records(sc, a, mode, rec != nullptr, st);
        if (rc_q)
            return rc_q;
        w.qrec = sc->qrec;
        w.obound = sc->q_obound;
    }
    w.list = list;
    w.rec = rec;
    w.min_tasks = split_at;
    w.list_count = sc->counters;
    w.cursor = sc->counters + 1;
    w.planes = planes;
    w.write_ids = write_ids;
    w.coop = env_int("RTC_NO_COOP", 0) == 0;
    w.spec = env_int("RTC_NO_SPEC", 0) == 0;
    w.spec_fails = env_int("RTC_SPEC_FAILS", 2);
    w.spec_retry = env_int("RTC_SPEC_RETRY", 0);
    w.tile = env_int("RTC_PATH_TILE", rtc::PATH_TILE);
    if (w.tile < 1 || w.tile > rtc::PATH_TILE)
        w.tile = rtc::PATH_TILE;
    if (rtc::mode_packed(mode) && (w.tile & 1))
        w.tile += 1; // pair records: a tile starts on an even triangle (PATH_TILE is even)
    const size_t resident = rtc::path_smem_bytes(sc->ntris, false, mode);
    const bool stream = env_int("RTC_FORCE_STREAM", 0) != 0 || resident > 96 * 1024;
    const size_t smem = rtc::path_smem_bytes(sc->ntris, stream, mode);
    int per_sm = 0;
    using PathKernel = void (*)(const rtc::RenderArgs, const rtc::PathWork);
    // [groups][stream][mode]: both group variants are launched, the device picks (see the kernel)
#define RTC_PATH_ROW(S_, G_)                                                                                       \
    {rtc::render_paths_kernel<S_, 0, G_>, rtc::render_paths_kernel<S_, 1, G_>, rtc::render_paths_kernel<S_, 2, G_>, \
     rtc::render_paths_kernel<S_, 3, G_>, rtc::render_paths_kernel<S_, 4, G_>}
    static const PathKernel kernels[2][2][5] = {{RTC_PATH_ROW(false, false), RTC_PATH_ROW(true, false)},
                                                {RTC_PATH_ROW(false, true), RTC_PATH_ROW(true, true)}};
#undef RTC_PATH_ROW
    const PathKernel kern = kernels[0][stream ? 1 : 0][mode], kern_groups = kernels[1][stream ? 1 : 0][mode];
    int rc_o = kernel_occupancy(sc->device, reinterpret_cast<const void *>(kern), rtc::PATH_THREADS, smem, &per_sm);
    if (rc_o)
        return rc_o;
    if (per_sm < 1)
        return fail(RTC_ERR_CUDA, "path kernel does not fit on an SM (smem %zu)", smem);
    long long tasks = (long long)a.out_w * a.out_h * planes; // upper bound when a list is used
    long long want = (tasks + rtc::PATH_THREADS - 1) / rtc::PATH_THREADS;
    int blocks = (int)std::min<long long>((long long)sc->nsm * per_sm, want > 0 ? want : 1);
    int per_sm_g = 0;
    if ((rc_o = kernel_occupancy(sc->device, reinterpret_cast<const void *>(kern_groups), rtc::PATH_THREADS, smem, &per_sm_g)))
        return rc_o;
    if (per_sm_g < 1)
        return fail(RTC_ERR_CUDA, "path kernel (group mode) does not fit on an SM (smem %zu)", smem);
    const int blocks_g = sc->nsm * std::min(per_sm_g, rtc::PATH_GROUP_CTAS);
    // RTC_GROUP_OVERSUB (percent, default 100): let the group mode count on more lanes than are resident -- groups then
    // take their pixels in more than one wave (A/B knob for frames split over 2-4 GPUs)
    w.group_lanes = (unsigned)((unsigned long long)blocks_g * rtc::PATH_THREADS * (unsigned)std::max(1, env_int("RTC_GROUP_OVERSUB", 100)) / 100u);
    if (split_at != 0)
        w.group_lanes = 0; // short lists are the pool form's: the serial form's group mode is not needed
    kern<<<blocks, rtc::PATH_THREADS, smem, st>>>(a, w);
    count_launch();
    if (split_at == 0) {
        kern_groups<<<blocks_g, rtc::PATH_THREADS, smem, st>>>(a, w);
        count_launch();
    }
    CU(cudaGetLastError());
    if (pool_deferred)
        return launch_pool(sc, a, list, rec, planes, mode, split_at, st);
    return RTC_OK;
}

rtc::PrimaryFrame make_primary_frame(const Camera *cam, const RtcDeviceJob *job, int R, bool block)
{
    rtc::PrimaryFrame fr;
    fr.origin = to3(cam->origin);
    fr.ex = to3(cam->ex);
    fr.ey = to3(cam->ey);
    fr.ez = to3(cam->ez);
    fr.fov = cam->fov;
    const int W = job->width, H = job->height;
    const double h2 = (double)(H / 2);
    const double dxlo = (0 - W / 2) / h2, dxhi = (W - 1 - W / 2) / h2, dylo = (0 - H / 2) / h2,
                 dyhi = (H - 1 - H / 2) / h2;
    fr.dxmax = fmax(fabs(dxlo), fabs(dxhi));
    fr.dymax = fmax(fabs(dylo), fabs(dyhi));
    fr.block = block ? 1 : 0;
    fr.hw = fr.hh = 0;
    if (block) {
        // half extents of an 8 x 4 pixel block in (dx, dy); inflated for the rounding of the block centre
        const int cs0 = job->subset.col_step > 0 ? job->subset.col_step : 1;
        const int bh = job->subset.band_h > 0 ? job->subset.band_h : 1;
        const int rs0 = bh == 1 ? (job->subset.row_step > 0 ? job->subset.row_step : 1) : 1;
        fr.hw = 0.5 * (rtc::BLK_W - 1) * cs0 / h2;
        fr.hh = 0.5 * (rtc::BLK_H - 1) * rs0 / h2;
        // blocks that stick out of the image have their centre up to one block beyond the last pixel
        fr.dxmax += 2.0 * fr.hw + 1.0 / h2;
        fr.dymax += 2.0 * fr.hh + 1.0 / h2;
        fr.hw = fr.hw * 1.001 + 4e-7 * (fr.dxmax + 1.0);
        fr.hh = fr.hh * 1.001 + 4e-7 * (fr.dymax + 1.0);
    }
    double dmax = 0;
    for (double dx : {dxlo, dxhi})
        for (double dy : {dylo, dyhi}) {
            double v[3] = {cam->ex.x * dx + cam->ey.x * dy + (double)cam->ez.x * cam->fov,
                           cam->ex.y * dx + cam->ey.y * dy + (double)cam->ez.y * cam->fov,
                           cam->ex.z * dx + cam->ey.z * dy + (double)cam->ez.z * cam->fov};
            dmax = fmax(dmax, sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]));
        }
    fr.Dmax = dmax * 1.001; // |D|^2 is convex in (dx, dy): the maximum over the image is at a corner
    int cs = job->subset.col_step > 0 ? job->subset.col_step : 1;
    fr.xstep = cs / h2;
    fr.max_steps = block ? 0 : R - 1;
    return fr;
}

int planes_of(const RtcDeviceJob *job)
{
    int G = job->nsplit > 1 ? job->nsplit : 1;
    int cnt = job->split_count > 0 ? job->split_count : G - job->split_first;
    return cnt;
}

} // namespace

extern "C" {

uint64_t rtc_launches_total(void) { return g_launches.load(); }

int rtc_ipc_export(int device, const void *d_ptr, unsigned char handle[RTC_IPC_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == RTC_IPC_HANDLE_BYTES, "CUDA IPC handles are 64 bytes");
    if (!d_ptr || !handle)
        return fail(RTC_ERR_ARG, "ipc_export: NULL argument");
    CU(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, const_cast<void *>(d_ptr)));
    memcpy(handle, &h, sizeof h);
    return RTC_OK;
}

int rtc_ipc_open(int device, const unsigned char handle[RTC_IPC_HANDLE_BYTES], void **d_ptr)
{
    if (!d_ptr || !handle)
        return fail(RTC_ERR_ARG, "ipc_open: NULL argument");
    CU(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    CU(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RTC_OK;
}

int rtc_ipc_close(int device, void *d_ptr)
{
    if (!d_ptr)
        return RTC_OK;
    CU(cudaSetDevice(device));
    CU(cudaIpcCloseMemHandle(d_ptr));
    return RTC_OK;
}

int rtc_enable_peer(int device, int peer)
{
    if (device == peer)
        return RTC_OK;
    int ok = 0;
    CU(cudaDeviceCanAccessPeer(&ok, device, peer));
    if (!ok)
        return fail(RTC_ERR_CUDA, "device %d cannot access device %d's memory", device, peer);
    CU(cudaSetDevice(device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
        return fail(RTC_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", device, peer, cudaGetErrorString(e));
    cudaGetLastError();
    return RTC_OK;
}

int rtc_render_launch_count(const RtcScene *sc, const RtcDeviceJob *job, int want_ids)
{
    (void)sc;
    (void)want_ids;
    if (!job || job->out_w <= 0 || job->out_h <= 0)
        return 0;
    return 5; // prepare_primary + block kernel + block-task kernel + the two path-kernel instantiations, one of
              // which returns at once (other modes launch 1-3)
}

static int render_device_impl(const RtcScene *sc, const Scene *sky, const Camera *cam, const RtcDeviceJob *job,
                              uint8_t *d_rgb, float *d_accum, int32_t *d_ids, uint64_t *d_stats, void *stream);

int rtc_render_device(const RtcScene *sc, const Scene *sky, const Camera *cam, const RtcDeviceJob *job,
                      uint8_t *d_rgb, float *d_accum, int32_t *d_ids, uint64_t *d_stats, void *stream)
{
    RtcScene *msc = const_cast<RtcScene *>(sc);
    if (msc)
        msc->sky_recorded = false;
    int rc = render_device_impl(sc, sky, cam, job, d_rgb, d_accum, d_ids, d_stats, stream);
    if (msc && msc->prog) {
        // whatever path the frame took, every chunk is final once the stream gets here
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        if (!rc) {
            if (!msc->sky_recorded)
                CU(cudaEventRecord(msc->sky_ev, st));
            CU(cudaMemsetAsync(msc->prog + 2 * msc->prog_chunks, 0x01, sizeof(unsigned) * msc->prog_chunks, st));
        }
        msc->prog = nullptr; // one-shot
    }
    return rc;
}

static int render_device_impl(const RtcScene *sc, const Scene *sky, const Camera *cam, const RtcDeviceJob *job,
                              uint8_t *d_rgb, float *d_accum, int32_t *d_ids, uint64_t *d_stats, void *stream)
{
    int rc = check_job(sc, sky, cam, job);
    if (rc)
        return rc;
    if (job->out_w == 0 || job->out_h == 0)
        return RTC_OK;
    int G = job->nsplit > 1 ? job->nsplit : 1;
    int planes = planes_of(job);
    if (job->split_first < 0 || planes <= 0 || job->split_first + planes > G)
        return fail(RTC_ERR_ARG, "render: sub-stream range [%d,+%d) outside nsplit %d", job->split_first, planes, G);
    if (G > 1 && !d_accum)
        return fail(RTC_ERR_ARG, "render: nsplit > 1 needs d_accum (planes of partial sums)");
    CU(cudaSetDevice(sc->device));
    {
        // a stale error of an earlier, unrelated runtime call must not be blamed on the launches below
        cudaError_t stale = cudaGetLastError();
        if (stale != cudaSuccess && getenv("RTC_TRACE"))
            fprintf(stderr, "rtc_b200: cleared stale CUDA error before render: %s\n", cudaGetErrorString(stale));
    }
    rtc::RenderArgs a = make_args(sc, sky, cam, job);
    a.rgb = d_rgb;
    a.accum = d_accum;
    a.ids = d_ids;
    a.stats = reinterpret_cast<unsigned long long *>(d_stats);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int R = fast_path_R(sc, cam, job);
    RtcScene *msc = const_cast<RtcScene *>(sc); // per-frame tables and work lists are scratch owned by the handle
    if (R > 0 || (job->flags & RTC_FLAG_PATHS_ONLY)) {
        const size_t nblocks = (size_t)((job->out_w + rtc::BLK_W - 1) / rtc::BLK_W + 8) *
                               (size_t)((job->out_h + rtc::BLK_H - 1) / rtc::BLK_H + 16) * (size_t)planes;
        if ((rc = ensure_scratch(msc, (size_t)job->out_w * job->out_h, nblocks)))
            return rc;
        CU(cudaMemsetAsync(msc->counters, 0, 8 * sizeof(unsigned), st));
    }
    const bool timed = (job->flags & RTC_FLAG_TIME_KERNELS) != 0 && R > 0;
    msc->kev_valid = false;
    if (timed) {
        for (auto &e : msc->kev)
            if (!e)
                CU(cudaEventCreate(&e));
        CU(cudaEventRecord(msc->kev[0], st));
    }
    if (R > 0) {
        if (sc->ntris > msc->cap_tab || !msc->tab) {
            CU(cudaFree(msc->tab));
            msc->tab = nullptr;
            CU(cudaMalloc(&msc->tab, sizeof(float4) * rtc::TAB_F4 * (size_t)(sc->ntris > 0 ? sc->ntris : 1)));
            CU(cudaMemsetAsync(msc->tab, 0, sizeof(float4) * rtc::TAB_F4 * (size_t)(sc->ntris > 0 ? sc->ntris : 1), st));
            msc->cap_tab = sc->ntris > 0 ? sc->ntris : 1;
        }
        const int nsph_eff = job->trianglesOnly ? 0 : sc->nsph;
        const int bh_ = job->subset.band_h > 0 ? job->subset.band_h : 1;
        const bool block = nsph_eff <= 32 && !(job->flags & RTC_FLAG_PER_PAIR_FILTER) &&
                           (bh_ == 1 || bh_ % rtc::BLK_H == 0) && env_int("RTC_SKY_BLOCK", 1) != 0;
        rtc::PrimaryFrame fr = make_primary_frame(cam, job, R, block);
        if (sc->ntris > 0) {
            rtc::prepare_primary_kernel<<<(sc->ntris + 127) / 128, 128, 0, st>>>(a.sc, fr, msc->tab);
            count_launch();
            CU(cudaGetLastError());
        }
        if (block && nsph_eff > 0) {
            if (nsph_eff > msc->cap_sphq) {
                CU(cudaFree(msc->sphq));
                msc->sphq = nullptr;
                CU(cudaMalloc(&msc->sphq, sizeof(float) * 8 * (size_t)nsph_eff));
                msc->cap_sphq = nsph_eff;
            }
            rtc::prepare_spheres_kernel<<<(nsph_eff + 31) / 32, 32, 0, st>>>(a.sc, fr, msc->sphq);
            count_launch();
            CU(cudaGetLastError());
        }
        rtc::SkyWork sw;
        sw.sphq = msc->sphq;
        sw.hw = (float)fr.hw;
        sw.hh = (float)fr.hh;
        sw.list = msc->list;
        sw.list_count = msc->counters;
        sw.rec = msc->rec;
        sw.btask = msc->btask;
        sw.btask_count = msc->counters + 2;
        sw.tile = env_int("RTC_SKY_TILE", rtc::SKY_TILE);
        if (sw.tile < 1 || sw.tile > rtc::SKY_TILE)
            sw.tile = rtc::SKY_TILE;
        sw.stream = (env_int("RTC_FORCE_STREAM", 0) != 0 || sc->ntris > rtc::SKY_TILE) ? 1 : 0;
        sw.segs = 1;
        if (block) {
            size_t bsmem = rtc::skyblock_smem_bytes(sc->ntris, sw.stream != 0);
            dim3 bgrid((job->out_w + 8 * rtc::BLK_W - 1) / (8 * rtc::BLK_W),
                       (job->out_h + 4 * rtc::BLK_H * rtc::SKYB_WARPS - 1) / (4 * rtc::BLK_H * rtc::SKYB_WARPS), planes);
            int sky_per_sm = 0;
            if ((rc = kernel_occupancy(sc->device, reinterpret_cast<const void *>(rtc::render_skyblock_kernel),
                                       rtc::SKYB_THREADS, bsmem, &sky_per_sm)))
                return rc;
            {
                // a short pixel grid (a frame split over several GPUs) leaves SMs idle while one warp re-sweeps its region
                // once per sample: split the samples over CTAs until the grid covers the machine twice
                const long long ctas = (long long)bgrid.x * bgrid.y * planes;
                const long long room = 2ll * msc->nsm * std::max(1, sky_per_sm);
                const int min_samples = std::max(1, job->spp / (job->nsplit > 1 ? job->nsplit : 1));
                int segs = (int)std::min<long long>(std::max<long long>(1, room / std::max<long long>(1, ctas)), 16);
                segs = std::max(1, std::min(segs, min_samples / 4));
                if (getenv("RTC_SKY_SEGS")) // tests / A-B: any count that leaves every segment of every sub-stream a sample
                    segs = std::max(1, std::min(std::min(env_int("RTC_SKY_SEGS", 1), 64), min_samples));
                sw.segs = segs;
                bgrid.z = (unsigned)(planes * segs);
            }
            if (msc->prog && job->nsplit <= 1 && d_rgb) {
                a.prog = msc->prog;
                a.prog_rows = msc->prog_rows;
                a.prog_chunks = msc->prog_chunks;
            }
            rtc::render_skyblock_kernel<<<bgrid, rtc::SKYB_THREADS, bsmem, st>>>(a, msc->tab, sw);
            count_launch();
            CU(cudaGetLastError());
            if (a.prog) { // sky pixels are final; hit pixels follow chunk by chunk from the path kernel
                CU(cudaEventRecord(msc->sky_ev, st));
                msc->sky_recorded = true;
            }
            if (timed)
                CU(cudaEventRecord(msc->kev[1], st));
            {
                // remaining samples of the blocks with candidate triangles: persistent warps over a work queue.
                // Independent of the path kernel (both only need the classification), so it is forked onto an
                // auxiliary stream and joined after the path kernel: the two latency-bound kernels share the SMs.
                if (!msc->aux) {
                    CU(cudaStreamCreateWithFlags(&msc->aux, cudaStreamNonBlocking));
                    CU(cudaEventCreateWithFlags(&msc->fork_ev, cudaEventDisableTiming));
                    CU(cudaEventCreateWithFlags(&msc->join_ev, cudaEventDisableTiming));
                }
                rtc::BlockTaskWork tw;
                tw.btask = msc->btask;
                tw.btask_count = msc->counters + 2;
                tw.cursor = msc->counters + 3;
                tw.tile = sw.tile;
                tw.stream = sw.stream;
                const size_t tsmem = rtc::sky_smem_bytes(sc->ntris, sw.stream != 0);
                int per_sm = 0;
                if ((rc = kernel_occupancy(sc->device, reinterpret_cast<const void *>(rtc::render_blocktask_kernel),
                                           rtc::BTASK_THREADS, tsmem, &per_sm)))
                    return rc;
                if (per_sm < 1)
                    return fail(RTC_ERR_CUDA, "block-task kernel does not fit on an SM");
                const bool overlap = env_int("RTC_NO_OVERLAP", 0) == 0;
                cudaStream_t ts = overlap ? msc->aux : st;
                if (overlap) {
                    CU(cudaEventRecord(msc->fork_ev, st));
                    CU(cudaStreamWaitEvent(msc->aux, msc->fork_ev, 0));
                }
                rtc::render_blocktask_kernel<<<msc->nsm * per_sm, rtc::BTASK_THREADS, tsmem, ts>>>(a, msc->tab, tw);
                count_launch();
                CU(cudaGetLastError());
                if (planes == 1 && env_int("RTC_NO_HINT", 0) == 0) { // this frame's list length, for the next frame's launch order (see launch_paths)
                    CU(cudaMemcpyAsync(msc->h_list_len, msc->counters, sizeof(unsigned), cudaMemcpyDeviceToHost, ts));
                    msc->list_len_seen = true;
                }
                if (overlap)
                    CU(cudaEventRecord(msc->join_ev, msc->aux));
                rc = launch_paths(msc, a, msc->list, msc->rec, planes, 0, job->flags, st);
                if (overlap)
                    CU(cudaStreamWaitEvent(st, msc->join_ev, 0));
            }
            if (timed && !rc) {
                CU(cudaEventRecord(msc->kev[2], st));
                msc->kev_valid = true;
            }
            return rc;
        }
        size_t smem = rtc::sky_smem_bytes(sc->ntris, sw.stream != 0);
        dim3 fgrid((job->out_w + rtc::FAST_LX * R - 1) / (rtc::FAST_LX * R),
                   (job->out_h + rtc::FAST_LY * rtc::FAST_WARPS - 1) / (rtc::FAST_LY * rtc::FAST_WARPS), planes);
#define LAUNCH_SKY(RR)                                                                                          \
    do {                                                                                                        \
        CU(cudaFuncSetAttribute(rtc::render_sky_kernel<RR>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                                (int)smem));                                                                    \
        rtc::render_sky_kernel<RR><<<fgrid, rtc::FAST_THREADS, smem, st>>>(a, msc->tab, sw);                     \
        count_launch(); \
    } while (0)
        if (R == 16)
            LAUNCH_SKY(16);
        else if (R == 8)
            LAUNCH_SKY(8);
        else if (R == 4)
            LAUNCH_SKY(4);
        else if (R == 2)
            LAUNCH_SKY(2);
        else
            LAUNCH_SKY(1);
#undef LAUNCH_SKY
        CU(cudaGetLastError());
        if (timed)
            CU(cudaEventRecord(msc->kev[1], st));
        rc = launch_paths(msc, a, msc->list, nullptr, planes, 0, job->flags, st);
        if (timed && !rc) {
            CU(cudaEventRecord(msc->kev[2], st));
            msc->kev_valid = true;
        }
        return rc;
    }
    if (job->flags & RTC_FLAG_PATHS_ONLY) {
        if (job->maxBounce < 1)
            return fail(RTC_ERR_ARG, "RTC_FLAG_PATHS_ONLY needs maxBounce >= 1");
        return launch_paths(msc, a, nullptr, nullptr, planes, 1, job->flags, st);
    }
    dim3 grid((job->out_w + rtc::CTA_W - 1) / rtc::CTA_W, (job->out_h + rtc::CTA_H - 1) / rtc::CTA_H, planes);
    if (job->flags & RTC_FLAG_COUNT_EXITS)
        rtc::render_kernel<true><<<grid, rtc::CTA_THREADS, 0, st>>>(a);
    else
        rtc::render_kernel<false><<<grid, rtc::CTA_THREADS, 0, st>>>(a);
    count_launch();
    CU(cudaGetLastError());
    return RTC_OK;
}

int rtc_scene_kernel_times(const RtcScene *sc, float *primary_ms, float *paths_ms)
{
    if (!sc || !sc->kev_valid)
        return fail(RTC_ERR_ARG, "kernel_times: the last render of this scene was not launched with RTC_FLAG_TIME_KERNELS");
    CU(cudaSetDevice(sc->device));
    CU(cudaEventSynchronize(sc->kev[2]));
    float a = 0, b = 0;
    CU(cudaEventElapsedTime(&a, sc->kev[0], sc->kev[1]));
    CU(cudaEventElapsedTime(&b, sc->kev[1], sc->kev[2]));
    if (primary_ms)
        *primary_ms = a;
    if (paths_ms)
        *paths_ms = b;
    return RTC_OK;
}

int rtc_sum_quantize_device(int device, const float *d_planes, int nplanes, size_t n, uint8_t *d_rgb, float *d_sum,
                            void *stream)
{
    if (!d_planes || nplanes <= 0)
        return fail(RTC_ERR_ARG, "sum_quantize: bad arguments");
    if (n == 0)
        return RTC_OK;
    CU(cudaSetDevice(device));
    rtc::sum_quantize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        d_planes, nplanes, n, d_rgb, d_sum);
    count_launch();
    CU(cudaGetLastError());
    return RTC_OK;
}

int rtc_quantize_device(int device, const float *d_accum, uint8_t *d_rgb, size_t n, void *stream)
{
    return rtc_sum_quantize_device(device, d_accum, 1, n, d_rgb, nullptr, stream);
}

} // extern "C"

// ---- render_cuda: the one-call drop-in for reference main.c:285-302 ------------------
//
// Single process, 1..8 devices.  Per device: upload + flatten the scene, render
// this device's share (tile split: interleaved 8-row bands; sample split:
// a contiguous range of RNG sub-streams), then gather on device 0 over NVLink
// with ncclAllGather (NCCL is dlopen'ed so a 1-GPU caller has no NCCL
// dependency), finish (un-interleave / ordered sum + quantise) and download.

#include <chrono>
#include <dlfcn.h>
#include <nccl.h>

namespace {

constexpr int BAND_H = 8;   // rows per interleaved band (two warp tiles)
constexpr int MAX_GPUS = 16;

struct NcclApi {
    void *h = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load()
    {
        if (h)
            return true;
        // RTC_NCCL_LIB: the NCCL the rest of the process uses (a Python caller points it at the copy bundled with
        // torch: a process that loads two libnccl.so.2 of different versions resolves the second one's users
        // against the first)
        const char *names[] = {getenv("RTC_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char *n : names)
            if (n && *n && (h = dlopen(n, RTLD_NOW | RTLD_LOCAL)))
                break;
        if (!h)
            return false;
#define SYM(field, name) field = reinterpret_cast<decltype(field)>(dlsym(h, name))
        SYM(CommInitAll, "ncclCommInitAll");
        SYM(CommDestroy, "ncclCommDestroy");
        SYM(AllGather, "ncclAllGather");
        SYM(GroupStart, "ncclGroupStart");
        SYM(GroupEnd, "ncclGroupEnd");
        SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
        return CommInitAll && CommDestroy && AllGather && GroupStart && GroupEnd && GetErrorString;
    }
};

struct DeviceSlot {
    RtcScene *scene = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr}; // upload begin, kernel begin, kernel end, gather end
    void *part = nullptr;   // this device's share (bytes or float planes); with NCCL it is a slice of `all`
    void *all = nullptr;    // gather buffer (G * part_bytes)
    size_t all_bytes = 0;
    int32_t *ids = nullptr;
    size_t ids_bytes = 0;
    unsigned long long *stats = nullptr;
    uint8_t *img = nullptr; // device 0: final W*H*3 image
    size_t img_bytes = 0;
    float *accum = nullptr; // tile split with nsplit > 1: local sub-stream planes
    size_t accum_bytes = 0;
    // single-GPU frames: the image goes to the host in row chunks, each as soon as its pixels are final
    cudaStream_t copy[4] = {nullptr, nullptr, nullptr, nullptr}; // chunk k goes out on copy[k % COPY_STREAMS]
    unsigned *prog = nullptr; // 3 * MAX_COPY_CHUNKS words, see RenderArgs::prog
    unsigned long long *h_stats = nullptr; // page-locked: the counters arrive with the frame, no extra blocking copy
};

constexpr int COPY_CHUNKS = 16;     // row chunks of a pipelined frame (RTC_COPY_CHUNKS, A/B knob)
constexpr int MAX_COPY_CHUNKS = 64;
constexpr int COPY_STREAMS = 4; // a chunk that waits for its last pixel only holds back every fourth chunk behind it
using WaitValue32Fn = int (*)(cudaStream_t, unsigned long long, unsigned, unsigned); // cuStreamWaitValue32
WaitValue32Fn wait_value32()
{
    static WaitValue32Fn fn = []() -> WaitValue32Fn {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<WaitValue32Fn>(p);
    }();
    return fn;
}

struct Context {
    std::mutex mu;
    DeviceSlot slot[MAX_GPUS];
    NcclApi nccl;
    ncclComm_t comms[MAX_GPUS];
    int ncomms = 0;
} g_ctx;

int slot_init(DeviceSlot &s, int dev)
{
    CU(cudaSetDevice(dev));
    if (!s.scene) {
        s.scene = new RtcScene;
        s.scene->device = dev;
    }
    if (!s.stream)
        CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    for (auto &e : s.ev)
        if (!e)
            CU(cudaEventCreate(&e));
    if (!s.stats)
        CU(cudaMalloc(&s.stats, sizeof(unsigned long long) * RTC_DEVICE_STATS_WORDS));
    if (!s.h_stats)
        CU(cudaMallocHost(&s.h_stats, sizeof(unsigned long long) * RTC_DEVICE_STATS_WORDS));
    return RTC_OK;
}

int ensure(void **p, size_t *have, size_t need)
{
    if (need <= *have && *p)
        return RTC_OK;
    cudaFree(*p);
    *p = nullptr;
    *have = 0;
    cudaError_t e = cudaMalloc(p, need ? need : 1);
    if (e != cudaSuccess)
        return fail(RTC_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", need, cudaGetErrorString(e));
    *have = need;
    return RTC_OK;
}

// compact band-interleaved parts [G][max_h][W] elements -> image rows, on device 0
template <typename T>
__global__ void uninterleave_kernel(const T *__restrict__ all, T *__restrict__ img, int W_elems, int H, int G,
                                    int band_h, size_t part_elems)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t total = (size_t)W_elems * H;
    if (i >= total)
        return;
    int y = (int)(i / W_elems), xe = (int)(i % W_elems);
    int band = y / band_h, g = band % G, j = (band / G) * band_h + y % band_h;
    img[i] = all[(size_t)g * part_elems + (size_t)j * W_elems + xe];
}

// Bands [band_lo, band_hi) of one part of a band split (the part's compact rows, band after band) to their rows of a HOST
// image: one strided device->host copy (+ one for a cut last band).
int download_band_range(int device, const void *d_part, void *h_img, size_t row_bytes, int height, int nparts, int band_h,
                        int part, int band_lo, int band_hi, cudaStream_t st)
{
    CU(cudaSetDevice(device));
    const size_t band_bytes = (size_t)band_h * row_bytes, stride = (size_t)nparts * band_bytes;
    // bands of this part: rows [b * nparts * band_h + part * band_h, + band_h) for b = 0, 1, ... while inside the image
    const int first = part * band_h, period = nparts * band_h;
    if (first >= height)
        return RTC_OK;
    const int nb = (height - first + period - 1) / period;                        // bands that start inside the image
    const int last_rows = std::min(band_h, height - (first + (nb - 1) * period)); // rows of the last one
    const int nfull = last_rows == band_h ? nb : nb - 1;
    const int lo = std::max(0, band_lo), hi = std::min(nb, band_hi);
    if (lo >= hi)
        return RTC_OK;
    char *dst = static_cast<char *>(h_img) + (size_t)first * row_bytes;
    const char *src = static_cast<const char *>(d_part);
    const int full_hi = std::min(hi, nfull);
    if (full_hi > lo)
        CU(cudaMemcpy2DAsync(dst + (size_t)lo * stride, stride, src + (size_t)lo * band_bytes, band_bytes, band_bytes,
                             (size_t)(full_hi - lo), cudaMemcpyDeviceToHost, st));
    if (hi > nfull && nfull >= lo) // the cut last band
        CU(cudaMemcpyAsync(dst + (size_t)nfull * stride, src + (size_t)nfull * band_bytes, (size_t)last_rows * row_bytes,
                           cudaMemcpyDeviceToHost, st));
    return RTC_OK;
}

} // namespace

namespace {
void stop_device_workers();
}

extern "C" void render_cuda_shutdown(void)
{
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    stop_device_workers();
    for (int i = 0; i < g_ctx.ncomms; ++i)
        g_ctx.nccl.CommDestroy(g_ctx.comms[i]);
    g_ctx.ncomms = 0;
    for (int d = 0; d < MAX_GPUS; ++d) {
        DeviceSlot &s = g_ctx.slot[d];
        if (!s.scene && !s.stream)
            continue;
        cudaSetDevice(d);
        rtc_scene_destroy(s.scene);
        if (s.stream)
            cudaStreamDestroy(s.stream);
        for (auto &e : s.ev)
            if (e)
                cudaEventDestroy(e);
        cudaFree(s.all);
        cudaFree(s.ids);
        cudaFree(s.stats);
        cudaFree(s.img);
        cudaFree(s.accum);
        cudaFree(s.prog);
        cudaFreeHost(s.h_stats);
        for (auto &cs : s.copy)
            if (cs)
                cudaStreamDestroy(cs);
        s = DeviceSlot();
    }
}

namespace {

#define NC(call)                                                                                   \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess)                                                                     \
            return fail(RTC_ERR_NCCL, "%s failed: %s", #call, g_ctx.nccl.GetErrorString(r_));      \
    } while (0)

int ensure_comms(int G)
{
    if (g_ctx.ncomms == G)
        return RTC_OK;
    if (!g_ctx.nccl.load())
        return fail(RTC_ERR_NCCL, "ngpus = %d needs libnccl.so.2 (dlopen failed: %s)", G, dlerror());
    for (int i = 0; i < g_ctx.ncomms; ++i)
        g_ctx.nccl.CommDestroy(g_ctx.comms[i]);
    g_ctx.ncomms = 0;
    int devs[MAX_GPUS];
    for (int i = 0; i < G; ++i)
        devs[i] = i;
    NC(g_ctx.nccl.CommInitAll(g_ctx.comms, G, devs));
    g_ctx.ncomms = G;
    return RTC_OK;
}

// One call of render_cuda: what every device's share looks like.
struct Call {
    const Scene *scene;
    const Camera *cam;
    uint8_t *out_rgb;
    const RenderParams *p;
    int W, H, G, S, base;
    bool samples;   // sample split (sub-stream planes gathered and summed on device 0)
    bool direct;    // tile split, G > 1: every device copies its own bands into the caller's image
    bool pipelined; // G == 1: row chunks go to the host as soon as their pixels are final
    size_t img_bytes, part_bytes;
    int chunk_rows, nchunks;
    RtcDeviceJob job[MAX_GPUS];
};

struct DevResult {
    int rc = RTC_OK;
    char err[512] = "";
    unsigned long long h[RTC_DEVICE_STATS_WORDS] = {0};
    float up_ms = 0, k_ms = 0;
    bool touched = false; // asynchronous work of this call may be in flight on the device's streams
};

int download_band_range(int device, const void *d_part, void *h_img, size_t row_bytes, int height, int nparts, int band_h,
                        int part, int band_lo, int band_hi, cudaStream_t st);

// Everything device g does for the frame, queued asynchronously: upload + flatten, render, and -- unless a collective
// follows (NCCL modes) -- the copy of its pixels into the caller's image.
int device_enqueue(const Call &c, int g, DevResult &r)
{
    const RenderParams *p = c.p;
    const int dev = c.base + g;
    DeviceSlot &s = g_ctx.slot[dev];
    int rc = slot_init(s, dev);
    if (rc)
        return rc;
    const bool gather = c.G > 1 && !c.direct;
    if ((rc = ensure(&s.all, &s.all_bytes, gather ? c.part_bytes * c.G : c.part_bytes)))
        return rc;
    s.part = static_cast<char *>(s.all) + (gather ? c.part_bytes * g : 0);
    const size_t npix = (size_t)c.job[g].out_w * c.job[g].out_h;
    const bool want_ids = p->out_primary_id && (!c.samples || g == 0);
    if (want_ids)
        if ((rc = ensure(reinterpret_cast<void **>(&s.ids), &s.ids_bytes, npix * sizeof(int32_t))))
            return rc;
    if (!c.samples && c.S > 1)
        if ((rc = ensure(reinterpret_cast<void **>(&s.accum), &s.accum_bytes, (size_t)c.S * npix * 3 * sizeof(float))))
            return rc;
    if (g == 0 && gather)
        if ((rc = ensure(reinterpret_cast<void **>(&s.img), &s.img_bytes, c.img_bytes)))
            return rc;
    r.touched = true;
    if (c.pipelined) {
        for (auto &cs : s.copy)
            if (!cs)
                CU(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
        if (!s.prog)
            CU(cudaMalloc(&s.prog, sizeof(unsigned) * 3 * MAX_COPY_CHUNKS));
        if (!s.scene->sky_ev)
            CU(cudaEventCreateWithFlags(&s.scene->sky_ev, cudaEventDisableTiming));
        CU(cudaMemsetAsync(s.prog, 0, sizeof(unsigned) * 3 * MAX_COPY_CHUNKS, s.stream));
        s.scene->prog = s.prog;
        s.scene->prog_rows = c.chunk_rows;
        s.scene->prog_chunks = c.nchunks;
    }
    CU(cudaEventRecord(s.ev[0], s.stream));
    CU(cudaMemsetAsync(s.stats, 0, sizeof(unsigned long long) * RTC_DEVICE_STATS_WORDS, s.stream));
    if ((rc = scene_upload(s.scene, p->tris, p->ntris, p->spheres, p->nspheres, s.stream)))
        return rc;
    CU(cudaEventRecord(s.ev[1], s.stream));
    int32_t *d_ids = want_ids ? s.ids : nullptr;
    uint64_t *d_stats = reinterpret_cast<uint64_t *>(s.stats);
    if (c.samples) {
        rc = rtc_render_device(s.scene, c.scene, c.cam, &c.job[g], nullptr, static_cast<float *>(s.part), d_ids, d_stats, s.stream);
    } else if (c.S > 1) {
        rc = rtc_render_device(s.scene, c.scene, c.cam, &c.job[g], nullptr, s.accum, d_ids, d_stats, s.stream);
        if (!rc)
            rc = rtc_sum_quantize_device(dev, s.accum, c.S, npix * 3, static_cast<uint8_t *>(s.part), nullptr, s.stream);
    } else {
        rc = rtc_render_device(s.scene, c.scene, c.cam, &c.job[g], static_cast<uint8_t *>(s.part), nullptr, d_ids, d_stats, s.stream);
    }
    if (rc)
        return rc;
    CU(cudaSetDevice(dev));
    CU(cudaEventRecord(s.ev[2], s.stream));
    // the counters travel with the frame (page-locked buffer): no blocking copy after the synchronisation
    CU(cudaMemcpyAsync(s.h_stats, s.stats, sizeof(unsigned long long) * RTC_DEVICE_STATS_WORDS, cudaMemcpyDeviceToHost, s.stream));
    if (c.pipelined) {
        // The copy streams follow the kernels: sky pixels are final after the classification kernel, the rows of a chunk
        // once the path kernel has finished its last primary-hit pixel (flag set on the device; rtc_render_device sets
        // every flag at the end of the frame whatever path it took).  Chunks go out round-robin on COPY_STREAMS streams so
        // that the sky-only chunks behind the geometry do not queue up behind the chunk that finishes last.
        const uint8_t *d_final = static_cast<const uint8_t *>(s.part);
        const int part_rows = c.job[g].out_h;
        for (auto &cs : s.copy)
            CU(cudaStreamWaitEvent(cs, s.scene->sky_ev, 0));
        for (int k = 0; k < c.nchunks; ++k) {
            const int r0 = k * c.chunk_rows;
            if (r0 >= part_rows)
                break;
            const int r1 = std::min(part_rows, r0 + c.chunk_rows);
            cudaStream_t cs = s.copy[k % COPY_STREAMS];
            if (wait_value32()(cs, (unsigned long long)(uintptr_t)(s.prog + 2 * c.nchunks + k), 1u, 0u /* GEQ */) != 0)
                return fail(RTC_ERR_CUDA, "cuStreamWaitValue32 failed");
            if (c.direct) { // this device's bands r0/8 .. of its compact part, to their rows of the caller's image
                if ((rc = download_band_range(dev, s.part, c.out_rgb, (size_t)c.W * 3, c.H, c.G, BAND_H, g, r0 / BAND_H,
                                              (r1 + BAND_H - 1) / BAND_H, cs)))
                    return rc;
            } else {
                CU(cudaMemcpyAsync(c.out_rgb + (size_t)r0 * c.W * 3, d_final + (size_t)r0 * c.W * 3, (size_t)(r1 - r0) * c.W * 3,
                                   cudaMemcpyDeviceToHost, cs));
            }
        }
    } else if (c.direct) {
        if ((rc = rtc_download_bands(dev, s.part, c.out_rgb, (size_t)c.W * 3, c.H, c.G, BAND_H, g, s.stream)))
            return rc;
    } else if (c.G == 1) {
        CU(cudaMemcpyAsync(c.out_rgb, s.part, c.img_bytes, cudaMemcpyDeviceToHost, s.stream));
    }
    return RTC_OK;
}

// wait for device g's queued work whatever happened before (a blocking call must not return with copies into the
// caller's buffers still in flight), and disarm the one-shot chunk flags
void device_drain(const Call &c, int g)
{
    DeviceSlot &s = g_ctx.slot[c.base + g];
    cudaSetDevice(c.base + g);
    if (s.stream)
        cudaStreamSynchronize(s.stream);
    for (auto &cs : s.copy)
        if (cs)
            cudaStreamSynchronize(cs);
    if (s.scene)
        s.scene->prog = nullptr;
    cudaGetLastError();
}

// after the frame: counters, event times, primary-hit ids of device g
int device_finish(const Call &c, int g, DevResult &r)
{
    const RenderParams *p = c.p;
    const int dev = c.base + g;
    DeviceSlot &s = g_ctx.slot[dev];
    CU(cudaSetDevice(dev));
    CU(cudaStreamSynchronize(s.stream));
    if (c.pipelined)
        for (auto &cs : s.copy)
            CU(cudaStreamSynchronize(cs));
    memcpy(r.h, s.h_stats, sizeof r.h);
    CU(cudaEventElapsedTime(&r.up_ms, s.ev[0], s.ev[1]));
    CU(cudaEventElapsedTime(&r.k_ms, s.ev[1], s.ev[2]));
    if (p->out_primary_id && (!c.samples || g == 0)) {
        const size_t npix = (size_t)c.job[g].out_w * c.job[g].out_h;
        if (c.G == 1 || c.samples) {
            CU(cudaMemcpy(p->out_primary_id, s.ids, npix * sizeof(int32_t), cudaMemcpyDeviceToHost));
        } else { // this device's bands of the id map (rows of different devices are disjoint)
            std::vector<int32_t> tmp(npix);
            CU(cudaMemcpy(tmp.data(), s.ids, npix * sizeof(int32_t), cudaMemcpyDeviceToHost));
            for (int j = 0; j < c.job[g].out_h; ++j) {
                const int y = c.job[g].subset.row_start + (j / BAND_H) * c.job[g].subset.row_step + j % BAND_H;
                memcpy(p->out_primary_id + (size_t)y * c.W, tmp.data() + (size_t)j * c.W, sizeof(int32_t) * c.W);
            }
        }
    }
    return RTC_OK;
}

// ---- one host thread per device (render_cuda with ngpus > 1): the devices' launch sequences, copies and
//      synchronisations run side by side instead of one after the other on the caller's thread ----
class DeviceWorkers {
  public:
    ~DeviceWorkers() { stop(); } // a process that never called render_cuda_shutdown() still joins its threads at exit
    void run(int n, const std::function<void(int)> &fn) // fn(1) .. fn(n - 1) on workers, fn(0) on the caller; waits for all
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            while ((int)threads_.size() < n - 1) {
                const int id = (int)threads_.size() + 1;
                threads_.emplace_back([this, id] { loop(id); });
            }
            fn_ = &fn;
            pending_ = n - 1;
            active_ = n;
            ++epoch_;
        }
        cv_.notify_all();
        fn(0);
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }
    void stop()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            quit_ = true;
        }
        cv_.notify_all();
        for (auto &t : threads_)
            t.join();
        threads_.clear();
        quit_ = false;
    }

  private:
    void loop(int id)
    {
        unsigned long long seen = 0;
        while (true) {
            const std::function<void(int)> *fn = nullptr;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return quit_ || (epoch_ != seen && id < active_); });
                if (quit_)
                    return;
                seen = epoch_;
                fn = fn_;
            }
            (*fn)(id);
            {
                std::lock_guard<std::mutex> lk(mu_);
                --pending_;
            }
            done_.notify_one();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_, done_;
    std::vector<std::thread> threads_;
    const std::function<void(int)> *fn_ = nullptr;
    unsigned long long epoch_ = 0;
    int pending_ = 0, active_ = 0;
    bool quit_ = false;
} g_workers;

void stop_device_workers() { g_workers.stop(); }

int render_locked(const Scene *scene, const Camera *cam, uint8_t *out_rgb, const RenderParams *p)
{
    Call c;
    c.scene = scene;
    c.cam = cam;
    c.out_rgb = out_rgb;
    c.p = p;
    c.W = p->width;
    c.H = p->height;
    c.G = p->ngpus > 1 ? p->ngpus : 1;
    c.base = 0;
    const unsigned long long launches0 = g_launches.load();
    const int W = c.W, H = c.H, G = c.G;
    const bool trace = getenv("RTC_TRACE") != nullptr;
    const auto t_entry = std::chrono::steady_clock::now();
    auto since = [&]() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_entry).count(); };
    int ndev = rtc_device_count();
    if (ndev <= 0)
        return fail(RTC_ERR_NO_GPU, "no CUDA device: librtc_b200 has no CPU fallback");
    if (G > ndev || G > MAX_GPUS)
        return fail(RTC_ERR_NO_GPU, "ngpus = %d but %d device(s) visible", G, ndev);
    if (G == 1) {
        const char *e = getenv("RTC_DEVICE");
        if (e)
            c.base = atoi(e);
        if (c.base < 0 || c.base >= ndev)
            return fail(RTC_ERR_NO_GPU, "RTC_DEVICE=%d but %d device(s) visible", c.base, ndev);
    }
    c.samples = G > 1 && p->split_mode == RTC_SPLIT_SAMPLES;
    c.S = p->nsplit > 1 ? p->nsplit : 1;
    if (c.samples) {
        // sub-stream seeds x + y*W + g*W*H are this library's own definition (not the reference's one stream per pixel):
        // the caller has to ask for them, so that the image never silently depends on the GPU count
        if (p->nsplit <= 1)
            return fail(RTC_ERR_ARG, "sample split: set nsplit (sub-streams per pixel, a multiple of ngpus = %d); only the tile "
                                     "split reproduces the reference's image", G);
        if (c.S % G)
            return fail(RTC_ERR_ARG, "sample split: nsplit (%d) must be a multiple of ngpus (%d)", c.S, G);
    }
    // Tile split: the caller's image is in HOST memory, so every device copies its own bands straight to their rows
    // (rtc_download_bands: one strided copy per device, each over its own PCIe link) -- no gather through device 0.
    // RTC_GATHER=1 keeps the frame's assembly on device 0 (NCCL all-gather over NVLink + un-interleave + one copy);
    // the sample split always needs it (ordered sum of the sub-stream planes).
    c.direct = G > 1 && !c.samples && env_int("RTC_GATHER", 0) == 0;
    const bool gather = G > 1 && !c.direct;
    if (gather) {
        int rc = ensure_comms(G);
        if (rc)
            return rc;
    }
    c.img_bytes = (size_t)W * H * 3;
    for (int g = 0; g < G; ++g) {
        RtcDeviceJob &j = c.job[g];
        memset(&j, 0, sizeof j);
        j.width = W; j.height = H; j.spp = p->spp; j.maxBounce = p->maxBounce; j.trianglesOnly = p->trianglesOnly;
        j.nsplit = c.S;
        j.flags = p->flags;
        if (c.samples) {
            j.split_first = g * (c.S / G);
            j.split_count = c.S / G;
            j.subset = RtcSubset{0, 1, 0, 1, 1};
        } else {
            j.split_first = 0;
            j.split_count = c.S;
            j.subset = G > 1 ? RtcSubset{g * BAND_H, G * BAND_H, 0, 1, BAND_H} : RtcSubset{0, 1, 0, 1, 1};
        }
        rtc_subset_size(W, H, &j.subset, &j.out_w, &j.out_h);
    }
    c.part_bytes = c.samples ? (size_t)(c.S / G) * c.img_bytes * sizeof(float) : (size_t)c.job[0].out_h * W * 3; // device 0 owns the most rows
    // one GPU, plain frame: the device->host copy runs chunk by chunk on its own stream, behind the kernels
    // (only frames whose copy takes long enough to be worth 2 x COPY_CHUNKS more runtime calls)
    // (G > 1, tile split: the same per device, over the rows of its compact part)
    const int part_rows = c.job[0].out_h;
    c.nchunks = std::max(1, std::min(env_int("RTC_COPY_CHUNKS", COPY_CHUNKS), MAX_COPY_CHUNKS));
    c.pipelined = (G == 1 || c.direct) && c.S == 1 && part_rows >= 8 * c.nchunks &&
                  c.part_bytes >= (size_t)env_int("RTC_PIPELINE_MIN_BYTES", 4 << 20) && wait_value32() != nullptr &&
                  getenv("RTC_NO_PIPELINE") == nullptr;
    c.chunk_rows = ((part_rows + c.nchunks - 1) / c.nchunks + 7) / 8 * 8; // whole 8 x 4 pixel blocks and whole bands

    DevResult res[MAX_GPUS];
    int rc = RTC_OK;
    double t_enq = 0;
    float gather_ms = 0;
    if (c.direct && env_int("RTC_NO_THREADS", 0) == 0) {
        // every device on its own host thread: enqueue, copy, synchronise, read the counters
        const std::function<void(int)> work = [&](int g) {
            DevResult &r = res[g];
            r.rc = device_enqueue(c, g, r);
            if (!r.rc)
                r.rc = device_finish(c, g, r);
            if (r.rc)
                snprintf(r.err, sizeof r.err, "%s", g_err);
        };
        g_workers.run(G, work);
        t_enq = since();
    } else {
        for (int g = 0; g < G && !rc; ++g)
            rc = device_enqueue(c, g, res[g]);
        DeviceSlot &s0 = g_ctx.slot[c.base];
        if (!rc && gather) {
            // gather over NVLink + finish on device 0, then one copy to the host
            rc = [&]() -> int {
                NC(g_ctx.nccl.GroupStart());
                for (int g = 0; g < G; ++g) {
                    DeviceSlot &s = g_ctx.slot[g];
                    NC(g_ctx.nccl.AllGather(s.part, s.all, c.part_bytes, ncclUint8, g_ctx.comms[g], s.stream));
                }
                NC(g_ctx.nccl.GroupEnd());
                CU(cudaSetDevice(c.base));
                if (c.samples) {
                    int r2 = rtc_sum_quantize_device(c.base, static_cast<const float *>(s0.all), c.S, c.img_bytes, s0.img, nullptr, s0.stream);
                    if (r2)
                        return r2;
                } else {
                    uninterleave_kernel<uint8_t><<<(unsigned)((c.img_bytes + 255) / 256), 256, 0, s0.stream>>>(
                        static_cast<const uint8_t *>(s0.all), s0.img, W * 3, H, G, BAND_H, c.part_bytes);
                    count_launch();
                    CU(cudaGetLastError());
                }
                CU(cudaEventRecord(s0.ev[3], s0.stream));
                CU(cudaMemcpyAsync(out_rgb, s0.img, c.img_bytes, cudaMemcpyDeviceToHost, s0.stream));
                return RTC_OK;
            }();
        }
        t_enq = since();
        for (int g = 0; g < G && !rc; ++g)
            rc = device_finish(c, g, res[g]);
        if (!rc && gather) {
            CU(cudaSetDevice(c.base));
            cudaEventElapsedTime(&gather_ms, s0.ev[2], s0.ev[3]);
        }
    }
    for (int g = 0; g < G; ++g)
        if (res[g].rc && !rc) {
            rc = res[g].rc;
            snprintf(g_err, sizeof g_err, "%s", res[g].err);
        }
    if (rc) {
        for (int g = 0; g < G; ++g)
            if (res[g].touched)
                device_drain(c, g);
        return rc;
    }
    const double t_sync = since();

    unsigned long long tot[RTC_DEVICE_STATS_WORDS] = {0};
    float up_ms = 0, k_ms = 0;
    for (int g = 0; g < G; ++g) {
        for (int i = 0; i < RTC_DEVICE_STATS_WORDS; ++i)
            tot[i] += res[g].h[i];
        up_ms = std::max(up_ms, res[g].up_ms);
        k_ms = std::max(k_ms, res[g].k_ms);
    }
    if (tot[rtc::ST_MISMATCH] >> 40)
        return fail(RTC_ERR_CUDA, "path kernel: scheduler watchdog tripped (%llu warp(s)); the image is incomplete",
                    tot[rtc::ST_MISMATCH] >> 40);
    if (p->stats) {
        RenderStats *st = p->stats;
        memset(st, 0, sizeof *st);
        st->rays = tot[rtc::ST_RAYS];
        st->tri_pairs = tot[rtc::ST_TRI_PAIRS];
        st->sphere_pairs = tot[rtc::ST_SPH_PAIRS];
        for (int e = 0; e < 6; ++e)
            st->tri_exit[e] = tot[rtc::ST_EXIT0 + e];
        st->hit_pixels = tot[rtc::ST_HIT_PIXELS];
        st->check_failures = tot[rtc::ST_MISMATCH];
        st->pool_samples_evaluated = tot[rtc::ST_POOL_EVAL];
        st->pool_shade_passes = tot[rtc::ST_POOL_SHADE_PASSES];
        st->pool_shade_lanes = tot[rtc::ST_POOL_SHADE_LANES];
        st->pool_sweep_passes = tot[rtc::ST_POOL_SWEEP_PASSES];
        st->pool_sweep_rays = tot[rtc::ST_POOL_SWEEP_RAYS];
        st->pool_iterations = tot[rtc::ST_POOL_ITERS];
        st->upload_ms = up_ms;
        st->kernel_ms = k_ms;
        st->gather_ms = gather_ms;
        st->download_ms = 0; // folded into the caller's wall clock; the copies are asynchronous to the kernels' events
        st->ngpus_used = G;
        st->kernels_launched = (int32_t)(g_launches.load() - launches0);
    }
    if (trace)
        fprintf(stderr, "rtc_b200: render_cuda host timeline: enqueued at %.0f us, synchronised at %.0f us, done at %.0f us\n",
                t_enq, t_sync, since());
    return RTC_OK;
}

} // namespace

extern "C" int render_cuda(const Scene *scene, const Camera *cam, uint8_t *out_rgb, const RenderParams *p)
{
    if (!scene || !cam || !out_rgb || !p)
        return fail(RTC_ERR_ARG, "render_cuda: NULL argument");
    if (p->width <= 0 || p->height < 2 || p->spp <= 0 || p->maxBounce < 0)
        return fail(RTC_ERR_ARG, "render_cuda: need width > 0, height >= 2, spp > 0, maxBounce >= 0");
    if (p->ntris < 0 || p->nspheres < 0 || (p->ntris > 0 && !p->tris) || (p->nspheres > 0 && !p->spheres))
        return fail(RTC_ERR_ARG, "render_cuda: triangle / sphere arrays inconsistent with their counts");
    if (p->split_mode != RTC_SPLIT_TILES && p->split_mode != RTC_SPLIT_SAMPLES)
        return fail(RTC_ERR_ARG, "render_cuda: unknown split_mode %d", p->split_mode);
    if (p->ngpus > 1 && p->split_mode == RTC_SPLIT_SAMPLES && (p->nsplit <= 1 || p->nsplit % p->ngpus))
        // sub-stream seeds x + y*W + g*W*H are this library's own definition (not the reference's one stream per pixel):
        // the caller has to ask for them, so that the image never silently depends on the GPU count
        return fail(RTC_ERR_ARG, "render_cuda: the sample split needs nsplit (sub-streams per pixel) set to a multiple of "
                                 "ngpus = %d; only the tile split reproduces the reference's image", p->ngpus);
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    return render_locked(scene, cam, out_rgb, p);
}

// ---- small device utilities exposed for multi-process drivers and the benchmark ------

namespace {

// dependent-free FFMA throughput probe: 8 independent chains per thread
__global__ void __launch_bounds__(256) ffma_probe_kernel(float *out, int iters, float a, float b)
{
    float r0 = threadIdx.x, r1 = r0 + 1, r2 = r0 + 2, r3 = r0 + 3, r4 = r0 + 4, r5 = r0 + 5, r6 = r0 + 6, r7 = r0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            r0 = fmaf(r0, a, b); r1 = fmaf(r1, a, b); r2 = fmaf(r2, a, b); r3 = fmaf(r3, a, b);
            r4 = fmaf(r4, a, b); r5 = fmaf(r5, a, b); r6 = fmaf(r6, a, b); r7 = fmaf(r7, a, b);
        }
    }
    float s = r0 + r1 + r2 + r3 + r4 + r5 + r6 + r7;
    if (s == 123.456f) // never true for the probe's inputs; keeps the chains alive
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

} // namespace

extern "C" {

// ---- test probes: the device versions of two reference functions on arrays (host buffers) ----
namespace {
__global__ void probe_random_direction_kernel(const uint32_t *states, int n, float *out, uint32_t *out_states)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float4 r = rtc::random_direction(states[i]);
    out[3 * (size_t)i] = r.x;
    out[3 * (size_t)i + 1] = r.y;
    out[3 * (size_t)i + 2] = r.z;
    out_states[i] = __float_as_uint(r.w);
}
__global__ void probe_environment_light_kernel(const float *dirs, int n, rtc::SkyParams sky, float *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const rtc::f3 e = rtc::environment_light(rtc::mk3(dirs[3 * (size_t)i], dirs[3 * (size_t)i + 1], dirs[3 * (size_t)i + 2]), sky);
    out[3 * (size_t)i] = e.x;
    out[3 * (size_t)i + 1] = e.y;
    out[3 * (size_t)i + 2] = e.z;
}
} // namespace

// Exhaustive self-check of the normal deviate's fast path (rtc_math.cuh): for the `count` generator states from `first`
// on, the two draws a deviate consumes, then the fast path against the exact sequence, bit for bit.
// out[0] = states where the fast path decided AND differs (must be 0), out[1] = states it declined (fall back), out[2] = states.
namespace {
__global__ void probe_normal_selfcheck_kernel(uint32_t first, unsigned long long count, unsigned long long *out)
{
    unsigned long long bad = 0, declined = 0, n = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < count;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        uint32_t st = first + (uint32_t)i;
        const float u1 = rtc::random_value(st), u2 = rtc::random_value(st);
        float v;
        const bool ok = rtc::nf::normal_fast(u1, u2, v);
        const float ref = rtc::normal_from_draws_exact(u1, u2);
        if (!ok)
            declined++;
        else if (__float_as_uint(v) != __float_as_uint(ref))
            bad++;
        n++;
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_xor_sync(0xffffffffu, bad, o);
        declined += __shfl_xor_sync(0xffffffffu, declined, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&out[0], bad);
        atomicAdd(&out[1], declined);
        atomicAdd(&out[2], n);
    }
}
} // namespace

int rtc_probe_normal_selfcheck(int device, uint32_t first_state, uint64_t count, uint64_t out[3])
{
    if (!out)
        return fail(RTC_ERR_ARG, "probe_normal_selfcheck: NULL");
    if (rtc_device_count() <= device || device < 0)
        return fail(RTC_ERR_NO_GPU, "probe_normal_selfcheck: device %d not visible", device);
    CU(cudaSetDevice(device));
    unsigned long long *d = nullptr;
    CU(cudaMalloc(&d, 3 * sizeof(unsigned long long)));
    CU(cudaMemset(d, 0, 3 * sizeof(unsigned long long)));
    int nsm = 0;
    CU(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device));
    probe_normal_selfcheck_kernel<<<nsm * 8, 256>>>(first_state, (unsigned long long)count, d);
    count_launch();
    CU(cudaGetLastError());
    unsigned long long h[3];
    CU(cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost));
    cudaFree(d);
    out[0] = h[0];
    out[1] = h[1];
    out[2] = h[2];
    return RTC_OK;
}

int rtc_probe_random_direction(int device, const uint32_t *states, int32_t n, float *out, uint32_t *out_states)
{
    if (n < 0 || (n > 0 && (!states || !out || !out_states)))
        return fail(RTC_ERR_ARG, "probe_random_direction: bad arguments");
    if (rtc_device_count() <= device || device < 0)
        return fail(RTC_ERR_NO_GPU, "probe_random_direction: device %d not visible", device);
    if (n == 0)
        return RTC_OK;
    CU(cudaSetDevice(device));
    uint32_t *d_s = nullptr, *d_os = nullptr;
    float *d_o = nullptr;
    CU(cudaMalloc(&d_s, sizeof(uint32_t) * (size_t)n));
    CU(cudaMalloc(&d_os, sizeof(uint32_t) * (size_t)n));
    CU(cudaMalloc(&d_o, sizeof(float) * 3 * (size_t)n));
    CU(cudaMemcpy(d_s, states, sizeof(uint32_t) * (size_t)n, cudaMemcpyHostToDevice));
    probe_random_direction_kernel<<<(n + 127) / 128, 128>>>(d_s, n, d_o, d_os);
    count_launch();
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, d_o, sizeof(float) * 3 * (size_t)n, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(out_states, d_os, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(d_s);
    cudaFree(d_os);
    cudaFree(d_o);
    return RTC_OK;
}

int rtc_probe_environment_light(int device, const Scene *sky, const float *dirs, int32_t n, float *out)
{
    if (!sky || n < 0 || (n > 0 && (!dirs || !out)))
        return fail(RTC_ERR_ARG, "probe_environment_light: bad arguments");
    if (rtc_device_count() <= device || device < 0)
        return fail(RTC_ERR_NO_GPU, "probe_environment_light: device %d not visible", device);
    if (n == 0)
        return RTC_OK;
    CU(cudaSetDevice(device));
    float *d_d = nullptr, *d_o = nullptr;
    CU(cudaMalloc(&d_d, sizeof(float) * 3 * (size_t)n));
    CU(cudaMalloc(&d_o, sizeof(float) * 3 * (size_t)n));
    CU(cudaMemcpy(d_d, dirs, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice));
    probe_environment_light_kernel<<<(n + 127) / 128, 128>>>(d_d, n, make_sky(sky), d_o);
    count_launch();
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, d_o, sizeof(float) * 3 * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(d_d);
    cudaFree(d_o);
    return RTC_OK;
}

// Measured FP32 FFMA peak of `device` (MEASURED_PEAKS.json has no FP32 entry): best of `reps` launches.
int rtc_fp32_peak_probe(int device, int reps, double *tflops, double *ms_best)
{
    if (!tflops)
        return fail(RTC_ERR_ARG, "fp32_peak_probe: NULL");
    if (rtc_device_count() <= device || device < 0)
        return fail(RTC_ERR_NO_GPU, "fp32_peak_probe: device %d not visible", device);
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    const int iters = 4096, threads = 256, blocks = prop.multiProcessorCount * 8 * 4;
    float *buf = nullptr;
    CU(cudaMalloc(&buf, sizeof(float) * (size_t)threads * blocks));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    double best = 1e30;
    for (int r = 0; r < (reps > 0 ? reps : 5) + 2; ++r) {
        CU(cudaEventRecord(e0));
        ffma_probe_kernel<<<blocks, threads>>>(buf, iters, 1.000001f, 0.5f);
        count_launch();
        CU(cudaEventRecord(e1));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        if (r >= 2 && ms < best)
            best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    double flops = 2.0 * 64.0 * iters * (double)threads * blocks;
    *tflops = flops / (best * 1e-3) / 1e12;
    if (ms_best)
        *ms_best = best;
    return RTC_OK;
}

// parts: [G][part_rows][row_elems] (rank g holds the 8-row bands g, g+G, ...) -> img [H][row_elems]
int rtc_uninterleave_device(int device, const void *d_parts, void *d_img, int elem_size, int row_elems, int height,
                            int nparts, int band_h, size_t part_elems, void *stream)
{
    if (!d_parts || !d_img || nparts <= 0 || band_h <= 0 || (elem_size != 1 && elem_size != 4))
        return fail(RTC_ERR_ARG, "uninterleave: bad arguments");
    CU(cudaSetDevice(device));
    size_t total = (size_t)row_elems * height;
    unsigned blocks = (unsigned)((total + 255) / 256);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (elem_size == 1)
        uninterleave_kernel<uint8_t><<<blocks, 256, 0, st>>>(static_cast<const uint8_t *>(d_parts),
                                                             static_cast<uint8_t *>(d_img), row_elems, height, nparts,
                                                             band_h, part_elems);
    else
        uninterleave_kernel<uint32_t><<<blocks, 256, 0, st>>>(static_cast<const uint32_t *>(d_parts),
                                                              static_cast<uint32_t *>(d_img), row_elems, height,
                                                              nparts, band_h, part_elems);
    count_launch();
    CU(cudaGetLastError());
    return RTC_OK;
}

// One part of a band split straight into the caller's HOST image: the part's compact rows (bands part,
// part + nparts, ... of band_h rows each) go to their image rows with one strided copy (+ one for a cut last band).
// With one process or thread per GPU every device writes its own rows over its own PCIe link -- no gather, no
// un-interleave and no full-frame copy through device 0 when the image is wanted in host memory anyway.
int rtc_download_bands(int device, const void *d_part, void *h_img, size_t row_bytes, int height, int nparts, int band_h,
                       int part, void *stream)
{
    if (!d_part || !h_img || row_bytes == 0 || height <= 0 || nparts <= 0 || band_h <= 0 || part < 0 || part >= nparts)
        return fail(RTC_ERR_ARG, "download_bands: bad arguments");
    return download_band_range(device, d_part, h_img, row_bytes, height, nparts, band_h, part, 0, 1 << 30,
                               static_cast<cudaStream_t>(stream));
}

} // extern "C"

// ---- rtc_render_paths_noalloc: the path kernel with its scratch in a static device array ---------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Declaration and contract:
// include/rtc_b200.h.  This block only ADDS an entry point: it launches the existing kernels
// (quick_records_kernel, render_paths_kernel) with pointers into the array below instead of the scene handle's
// cudaMalloc'ed scratch, and changes no other code path.

namespace {

constexpr size_t NOALLOC_HEAD_BYTES = 256; // work cursor + counters of the path kernel; cleared by every call
constexpr size_t NOALLOC_QREC_F4 = (size_t)rtc::QPAIR_F4 * (RTC_NOALLOC_MAX_TRIS / 2 + 1); // pair records, packed layout
static_assert(NOALLOC_HEAD_BYTES + NOALLOC_QREC_F4 * sizeof(float4) == RTC_NOALLOC_SCRATCH_BYTES,
              "RTC_NOALLOC_SCRATCH_BYTES of rtc_b200.h is out of date");

// [cursor words (256 B)][quick records of up to RTC_NOALLOC_MAX_TRIS triangles].  One per GPU that loads the library,
// held until exit, shared by every call: it keeps the LAST call's contents.
static __device__ __align__(256) unsigned char g_noalloc_scratch[RTC_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, NOT a device address: the address a kernel
// may be given comes from cudaGetSymbolAddress, asked once per device with that device current.
int noalloc_scratch_address(int device, unsigned char **out)
{
    static unsigned char *addr[MAX_GPUS] = {nullptr};
    static std::mutex mu;
    if (device < 0 || device >= MAX_GPUS)
        return fail(RTC_ERR_ARG, "render_paths_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[device]) {
        CU(cudaSetDevice(device));
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p)
            return fail(RTC_ERR_CUDA, "render_paths_noalloc: cudaGetSymbolAddress returned a null address");
        addr[device] = static_cast<unsigned char *>(p);
    }
    *out = addr[device];
    return RTC_OK;
}

} // namespace

extern "C" int rtc_render_paths_noalloc(const RtcScene *sc, const Scene *sky, const Camera *cam, const RtcDeviceJob *job,
                                        uint8_t *d_rgb, int32_t *d_ids, uint64_t *d_stats, void *stream)
{
    int rc = check_job(sc, sky, cam, job);
    if (rc)
        return rc;
    if (!d_rgb)
        return fail(RTC_ERR_ARG, "render_paths_noalloc: d_rgb is required");
    if (sc->ntris > RTC_NOALLOC_MAX_TRIS)
        return fail(RTC_ERR_ARG, "render_paths_noalloc: the scene has %d triangles, the static scratch holds the records of at "
                                 "most RTC_NOALLOC_MAX_TRIS = %d (use rtc_render_device)", sc->ntris, RTC_NOALLOC_MAX_TRIS);
    if (job->nsplit > 1)
        return fail(RTC_ERR_ARG, "render_paths_noalloc: nsplit = %d; only the reference's one RNG stream per pixel", job->nsplit);
    if (job->maxBounce < 1)
        return fail(RTC_ERR_ARG, "render_paths_noalloc: needs maxBounce >= 1");
    if (job->out_w == 0 || job->out_h == 0)
        return RTC_OK;
    CU(cudaSetDevice(sc->device));
    unsigned char *scratch = nullptr;
    if ((rc = noalloc_scratch_address(sc->device, &scratch)))
        return rc;
    cudaGetLastError(); // a stale error of an earlier, unrelated call must not be blamed on the launches below
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    rtc::RenderArgs a = make_args(sc, sky, cam, job);
    a.rgb = d_rgb;
    a.ids = d_ids;
    a.stats = reinterpret_cast<unsigned long long *>(d_stats);

    // what must start at zero: the work cursor (the array keeps the previous call's value)
    CU(cudaMemsetAsync(scratch, 0, NOALLOC_HEAD_BYTES, st));
    // quick records: packed form when the stored normals follow the winding (as rtc_render_device decides), else the
    // reference-arithmetic first phase, which needs no records.  Rebuilt by EVERY call: the array is shared by all
    // scenes and calls, whatever it holds belongs to the last one.
    const int mode = sc->quick_ok ? 3 : 0;
    rtc::PathWork w;
    memset(&w, 0, sizeof w);
    if (mode != 0) {
        float need = std::fmax(sc->bound_abs, std::fmax(std::fabs(a.origin.x), std::fmax(std::fabs(a.origin.y), std::fabs(a.origin.z))));
        need = (need * 1.001f + 1e-6f) * 1.25f; // the primary rays start at the camera: it counts for the origin bound
        float4 *qrec = reinterpret_cast<float4 *>(scratch + NOALLOC_HEAD_BYTES);
        if (sc->ntris > 0) {
            rtc::quick_records_kernel<<<(sc->ntris + 127) / 128, 128, 0, st>>>(sc->geo, sc->ntris, need, qrec, rtc::mode_layout(mode));
            count_launch();
            CU(cudaGetLastError());
        }
        w.qrec = qrec;
        w.obound = need;
    }
    w.list = nullptr; // every pixel of the subset; the kernel sweeps the primary rays itself
    w.rec = nullptr;
    w.list_count = reinterpret_cast<unsigned *>(scratch);
    w.cursor = reinterpret_cast<unsigned *>(scratch) + 1;
    w.planes = 1;
    w.write_ids = 1;
    w.tile = rtc::PATH_TILE;
    w.coop = 1;
    w.group_lanes = 0; // the plain (GROUPS = false) instantiation
    w.spec = 0;
    w.spec_fails = 2;
    w.spec_retry = 0;
    w.min_tasks = 0;
    const bool streamed = rtc::path_smem_bytes(sc->ntris, false, mode) > 96 * 1024; // tiles through shared memory (TMA)
    const size_t smem = rtc::path_smem_bytes(sc->ntris, streamed, mode);
    using PathKernel = void (*)(const rtc::RenderArgs, const rtc::PathWork);
    const PathKernel kern = mode == 3 ? (streamed ? rtc::render_paths_kernel<true, 3, false> : rtc::render_paths_kernel<false, 3, false>)
                                      : (streamed ? rtc::render_paths_kernel<true, 0, false> : rtc::render_paths_kernel<false, 0, false>);
    int per_sm = 0, nsm = 0;
    if ((rc = kernel_occupancy(sc->device, reinterpret_cast<const void *>(kern), rtc::PATH_THREADS, smem, &per_sm))) // no allocation
        return rc;
    if (per_sm < 1)
        return fail(RTC_ERR_CUDA, "render_paths_noalloc: path kernel does not fit on an SM (smem %zu)", smem);
    CU(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, sc->device));
    const long long tasks = (long long)a.out_w * a.out_h;
    const long long want = (tasks + rtc::PATH_THREADS - 1) / rtc::PATH_THREADS;
    const int blocks = (int)std::min<long long>((long long)nsm * per_sm, want > 0 ? want : 1);
    kern<<<blocks, rtc::PATH_THREADS, smem, st>>>(a, w);
    count_launch();
    CU(cudaGetLastError());
    return RTC_OK;
}
