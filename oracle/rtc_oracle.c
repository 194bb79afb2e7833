/*
 * TEST INFRASTRUCTURE -- not product code, never shipped, never on the
 * product path (the product fails loudly without its CUDA library).
 *
 * rtc_oracle.c: a plain-C, single-threaded RESTATEMENT of the reference's
 * per-pixel render loop, written from the reference's behaviour (each function
 * cites the reference file:line it follows; paths are into /root/reference).
 * It exists to (1) check the CUDA path on boxes where /root/reference is
 * absent, (2) count rays / ray-triangle exits, which the pristine reference
 * cannot report, and (3) serve as the "port" CPU baseline when oracle/_ref is
 * not built.
 *
 * PARITY PINNING: the reference ships no tests or golden vectors for this
 * path.  This restatement is pinned instead against the
 * reference itself: tests/test_oracle_golden.py compares it bit-for-bit
 * with oracle/_ref (the reference's own functions, compiled from
 * /root/reference) and tests/golden/ holds the vectors generated that way by
 * tests/golden/make_golden.py.
 *
 * Floating-point contract: every `float op float` below is one IEEE binary32
 * operation, never contracted (build: -ffp-contract=off, no -march, no
 * -ffast-math).  Where the reference computes in double (it mixes `1.`, `2.`,
 * sqrt(), log(), cos() into float code) the same widening is written out.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "oracle_job.h"

typedef OracleVec3 v3;

typedef struct { v3 color; float emission, smoothness; } OMat;         /* raytracing.h:25-30 */
typedef struct { v3 a, b, c, n; OMat mat; } OTri;                      /* raytracing.h:41-45 */
typedef struct { v3 pos; float r; OMat mat; } OSph;                    /* raytracing.h:34-39 */

_Static_assert(sizeof(OTri) == 68, "Triangle is 68 bytes");
_Static_assert(sizeof(OSph) == 36, "Sphere is 36 bytes");

static OTri *g_tris;
static int g_ntris;
static OSph *g_sph;
static int g_nsph;
static uint32_t g_rng;          /* scene.h:35 */
static OracleStats g_stats;

/* ---- moremath.c:7-87 ---------------------------------------------------- */

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }       /* :56 */
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }       /* :62 */
static inline v3 scale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }        /* :68 */
static inline v3 mul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }       /* :74 */
static inline float dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }    /* :34, left to right */
static inline v3 cross3(v3 u, v3 v)                                                   /* :44 */
{
    return V(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
/* :7-18  length() takes sqrt in double of a float sum and returns float;
   normalized() then forms 1./len in double and narrows it to float. */
static inline v3 unit(v3 v)
{
    float len = (float)sqrt((double)(v.x * v.x + v.y * v.y + v.z * v.z));
    float inv = (float)(1. / (double)len);
    return V(v.x * inv, v.y * inv, v.z * inv);
}
static inline float clamp01(float x) { return x < 0 ? 0 : (x > 1 ? 1 : x); }          /* :39 */
/* :50-54  float clamp of a float quotient; x*x in float; the cubic's second
   factor and the product in double; narrowed on return. */
static inline float smooth(float lo, float hi, float x)
{
    x = clamp01((x - lo) / (hi - lo));
    return (float)((double)(x * x) * (3. - 2. * (double)x));
}
/* :80-83  dir - normal * (float)(2. * dot) */
static inline v3 reflect3(v3 d, v3 n) { return sub(d, scale(n, (float)(2. * (double)dot3(d, n)))); }
/* :85-88  a*(1-t) + b*t */
static inline v3 mix(v3 a, v3 b, float t) { return add(scale(a, 1 - t), scale(b, t)); }

/* ---- moremath.c:90-109 RNG --------------------------------------------- */

static inline float rnd(void) /* :90-96 */
{
    g_rng = g_rng * 747796405u + 2891336453u;
    uint32_t r = ((g_rng >> ((g_rng >> 28) + 4)) ^ g_rng) * 277803737u;
    r = (r >> 22) ^ r;
    g_stats.rng_draws++;
    return (float)((double)r / 4294967295.0);
}
static inline float rnd_normal(void) /* :98-103, first draw is the angle */
{
    float theta = (float)(2 * 3.14159265 * (double)rnd());
    float rho = (float)sqrt(-2 * log((double)rnd()));
    return (float)((double)rho * cos((double)theta));
}
static inline v3 rnd_dir(void) /* :105-109, x then y then z */
{
    float x = rnd_normal();
    float y = rnd_normal();
    float z = rnd_normal();
    return unit(V(x, y, z));
}

/* ---- raytracing.c:151-160 sky ------------------------------------------ */

static v3 sky(v3 dir, const OracleScene *s)
{
    g_stats.sky_evals++;
    float t = powf(smooth(0, .74f, -dir.y), .35f);
    v3 grad = mix(s->horizon, s->zenith, t);
    float sun = powf((float)fmax(0, (double)dot3(dir, s->sun)), s->sunFocus) * s->sunIntensity;
    float g2s = smooth(-0.01f, 0, -dir.y);
    float mask = dir.y < 0;
    v3 sunv = V(sun * mask, sun * mask, sun * mask);
    return add(mix(s->ground, grad, g2s), sunv);
}

/* ---- raytracing.c:162-184 sphere --------------------------------------- */

static int hit_sphere(v3 pos, v3 dir, v3 centre, float radius, float *dst, v3 *normal)
{
    v3 off = sub(pos, centre);
    float b = dot3(off, dir);
    float c = dot3(off, off) - radius * radius;
    float delta = b * b - c;
    if (delta < 0)
        return 0;
    delta = (float)sqrt((double)delta);
    float d = -b - delta;
    if ((double)d < 0.001)
        d = -b + delta;
    if ((double)d < 0.001)
        return 0;
    *dst = d;
    v3 hp = add(pos, scale(dir, d));
    *normal = unit(sub(hp, centre));
    return 1;
}

/* ---- raytracing.c:186-214 triangle ------------------------------------- */

static int hit_triangle(v3 pos, v3 dir, const OTri *t, float *dst)
{
    if (dot3(dir, t->n) >= 0) { g_stats.tri_exit[0]++; return 0; }
    v3 ab = sub(t->b, t->a);
    v3 ac = sub(t->c, t->a);
    v3 h = cross3(dir, ac);
    float det = dot3(ab, h);
    if (-0.001 < (double)det && (double)det < 0.001) { g_stats.tri_exit[1]++; return 0; }
    float inv = (float)(1. / (double)det);
    v3 s = sub(pos, t->a);
    float u = dot3(s, h) * inv;
    if (u < 0 || u > 1) { g_stats.tri_exit[2]++; return 0; }
    v3 q = cross3(s, ab);
    float v = dot3(dir, q) * inv;
    if (v < 0 || u + v > 1) { g_stats.tri_exit[3]++; return 0; }
    float d = dot3(ac, q) * inv;
    if ((double)d < 0.001) { g_stats.tri_exit[4]++; return 0; }
    g_stats.tri_exit[5]++;
    *dst = d;
    return 1;
}

/* ---- raytracing.c:216-240 closest hit ---------------------------------- */

typedef struct { int hit; int32_t id; float dst; v3 point, normal; OMat mat; } OHit;

static OHit closest(v3 pos, v3 dir, int trianglesOnly)
{
    OHit best;
    memset(&best, 0, sizeof best);
    best.dst = 999999;
    best.id = -1;
    g_stats.rays++;
    if (trianglesOnly == 0)
        for (int i = 0; i < g_nsph; ++i) {
            float d;
            v3 n;
            g_stats.sphere_tests++;
            if (hit_sphere(pos, dir, g_sph[i].pos, g_sph[i].r, &d, &n) && d < best.dst) {
                best.hit = 1; best.dst = d; best.normal = n; best.mat = g_sph[i].mat; best.id = -2 - i;
            }
        }
    for (int i = 0; i < g_ntris; ++i) {
        float d;
        if (hit_triangle(pos, dir, &g_tris[i], &d) && d < best.dst) {
            best.hit = 1; best.dst = d; best.normal = g_tris[i].n; best.mat = g_tris[i].mat; best.id = i;
        }
    }
    best.point = add(pos, scale(dir, best.dst));
    return best;
}

/* ---- raytracing.c:262-296 one path ------------------------------------- */

static v3 trace_path(v3 pos, v3 dir, int trianglesOnly, int maxBounce, const OracleScene *s)
{
    v3 light = V(0, 0, 0), tint = V(1, 1, 1);
    for (int i = 0; i < maxBounce; ++i) {
        OHit h = closest(pos, dir, trianglesOnly);
        if (!h.hit) {
            light = add(light, mul(sky(dir, s), tint));
            break;
        }
        v3 diffuse = unit(add(h.normal, rnd_dir()));
        v3 specular = reflect3(dir, h.normal);
        dir = mix(diffuse, specular, h.mat.smoothness); /* not renormalised, :276 */
        pos = h.point;
        v3 emitted = scale(h.mat.color, h.mat.emission);
        light = add(light, mul(emitted, tint));
        tint = mul(tint, h.mat.color);
        float p = (float)fmax(fmax((double)tint.x, (double)tint.y), (double)tint.z);
        if (p < rnd())
            break;
        tint = scale(tint, (float)(1.0 / (double)p));
    }
    return light;
}

/* ---- moremath.c:25-30 quantise ----------------------------------------- */

static inline uint8_t quant(float f)
{
    if (f < 0)
        return 0;
    if (f >= 1)
        return 255;
    if (f != f)
        return 0; /* NaN: x86 cvttss2si gives 0x80000000, low byte 0 */
    return (uint8_t)(f * 255.f);
}

/* ---- exported ----------------------------------------------------------- */

void oracle_set_triangles(const void *tris, int n)
{
    free(g_tris);
    g_tris = malloc(sizeof(OTri) * (size_t)(n > 0 ? n : 1));
    memcpy(g_tris, tris, sizeof(OTri) * (size_t)n);
    g_ntris = n;
}

void oracle_set_spheres(const void *sph, int n)
{
    free(g_sph);
    g_sph = malloc(sizeof(OSph) * (size_t)(n > 0 ? n : 1));
    memcpy(g_sph, sph, sizeof(OSph) * (size_t)n);
    g_nsph = n;
}

/* main.c:247 and :252-255 */
void oracle_normalize(const float in[3], float out[3])
{
    v3 v = unit(V(in[0], in[1], in[2]));
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
}

void oracle_camera_look_at(const float origin[3], const float lookingAt[3], float fov, OracleCamera *cam)
{
    v3 o = V(origin[0], origin[1], origin[2]);
    v3 ez = unit(sub(V(lookingAt[0], lookingAt[1], lookingAt[2]), o));
    v3 ex = unit(cross3(ez, V(0, -1, 0)));
    v3 ey = unit(cross3(ez, ex));
    cam->origin = o; cam->ex = ex; cam->ey = ey; cam->ez = ez; cam->fov = fov;
}

/* main.c:84-101, with spp / size / pixel subset / nsplit as parameters */
void oracle_render(const OracleJob *job, uint8_t *rgb, int32_t *ids, OracleStats *stats)
{
    const int W = job->width, H = job->height, spp = job->spp;
    const int G = job->nsplit > 1 ? job->nsplit : 1;
    const int rstep = job->row_step > 0 ? job->row_step : 1;
    const int cstep = job->col_step > 0 ? job->col_step : 1;
    const float w = (float)(1. / spp);
    memset(&g_stats, 0, sizeof g_stats);

    for (int y = job->row_start; y < H; y += rstep)
        for (int x = job->col_start; x < W; x += cstep) {
            float dx = (x - W / 2) / (float)(H / 2);
            float dy = (y - (H / 2)) / (float)(H / 2);
            v3 dir = add(add(scale(job->cam.ex, dx), scale(job->cam.ey, dy)), scale(job->cam.ez, job->cam.fov));
            dir = unit(dir);
            if (ids) {
                OracleStats keep = g_stats; /* the id probe is not a counted ray */
                OHit h = closest(job->cam.origin, dir, job->trianglesOnly);
                g_stats = keep;
                ids[x + y * W] = h.id;
                g_stats.hit_pixels += h.hit;
            }
            v3 total = V(0, 0, 0);
            for (int g = 0; g < G; ++g) {
                g_rng = (uint32_t)(x + y * W) + (uint32_t)g * (uint32_t)(W * H);
                v3 acc = V(0, 0, 0);
                int s0 = (int)((long long)g * spp / G), s1 = (int)((long long)(g + 1) * spp / G);
                for (int i = s0; i < s1; ++i)
                    acc = add(acc, scale(trace_path(job->cam.origin, dir, job->trianglesOnly, job->maxBounce, &job->scene), w));
                total = G == 1 ? acc : add(total, acc);
            }
            if (rgb) {
                uint8_t *p = rgb + 3 * (size_t)(x + y * W);
                p[0] = quant(total.x); p[1] = quant(total.y); p[2] = quant(total.z);
            }
            g_stats.pixels++;
        }
    if (stats)
        *stats = g_stats;
}

/* probes for unit tests */
float oracle_random_value(uint32_t *state)
{
    g_rng = *state;
    float v = rnd();
    *state = g_rng;
    return v;
}

void oracle_random_direction(uint32_t *state, float out[3])
{
    g_rng = *state;
    v3 v = rnd_dir();
    *state = g_rng;
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
}

/* the same for n independent states (test probe at scale) */
void oracle_random_direction_batch(const uint32_t *states, int n, float *out, uint32_t *out_states)
{
    for (int i = 0; i < n; ++i) {
        uint32_t st = states[i];
        oracle_random_direction(&st, out + 3 * (size_t)i);
        out_states[i] = st;
    }
}

void oracle_environment_light(const float pos[3], const float dir[3], const OracleScene *s, float out[3])
{
    (void)pos;
    v3 v = sky(V(dir[0], dir[1], dir[2]), s);
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
}

/* n directions at once (test probe at scale) */
void oracle_environment_light_batch(const float *dirs, int n, const OracleScene *s, float *out)
{
    const float pos[3] = {0.f, 0.f, 0.f};
    for (int i = 0; i < n; ++i)
        oracle_environment_light(pos, dirs + 3 * (size_t)i, s, out + 3 * (size_t)i);
}

int oracle_ray_triangle(const float pos[3], const float dir[3], int i, float *dst)
{
    *dst = 0;
    return hit_triangle(V(pos[0], pos[1], pos[2]), V(dir[0], dir[1], dir[2]), &g_tris[i], dst);
}

int oracle_ray_sphere(const float pos[3], const float dir[3], int k, float *dst, float normal[3])
{
    v3 n = V(0, 0, 0);
    *dst = 0;
    int hit = hit_sphere(V(pos[0], pos[1], pos[2]), V(dir[0], dir[1], dir[2]), g_sph[k].pos, g_sph[k].r, dst, &n);
    normal[0] = n.x; normal[1] = n.y; normal[2] = n.z;
    return hit;
}

uint8_t oracle_float_to_uint(float f) { return quant(f); }

/* FNV-1a-64 as used by the known-answer table of SURVEY.md section 8c */
uint64_t oracle_fnv1a64_bytes(const uint8_t *p, uint64_t n)
{
    uint64_t h = 1469598103934665603ull;
    for (uint64_t i = 0; i < n; ++i)
        h = (h ^ p[i]) * 1099511628211ull;
    return h;
}

/* id hash: xor each whole uint32 word, then multiply */
uint64_t oracle_fnv1a64_words(const uint32_t *p, uint64_t n)
{
    uint64_t h = 1469598103934665603ull;
    for (uint64_t i = 0; i < n; ++i)
        h = (h ^ p[i]) * 1099511628211ull;
    return h;
}
