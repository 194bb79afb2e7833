/*
 * TEST INFRASTRUCTURE -- not product code, never shipped, never measured as
 * the product.
 *
 * Single-translation-unit build of the REFERENCE's own functions, compiled
 * from the sources where they lie (-I/root/reference; nothing is copied into
 * this repository).  `#pragma once` in the reference headers collapses the
 * header-static globals of scene.h (triangles, triangleCount, rngState) into
 * one copy, so the per-pixel seed `rngState = x + y*width` (main.c:95) is live
 * and a single-threaded pixel loop is bit-reproducible.
 *
 * This file holds NO copy of reference logic except the pixel loop of
 * main.c:84-101 (ray generation, seed, spp accumulation, quantise), which has
 * to be restated because the reference hard-codes spp (scene.h:26) and keeps
 * the loop inside a pthread entry point.  Everything called from that loop --
 * calcColor, rayTriangle, raySphere, normalized, ... -- is the reference's.
 *
 * Built by oracle/Makefile into oracle/_ref/libref_oracle.so.
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <fcntl.h>

#include "moremath.c"   /* /root/reference/moremath.c   */
#include "raytracing.c" /* /root/reference/raytracing.c */
#include "objloader.c"  /* /root/reference/objloader.c  */

#include "oracle_job.h"

_Static_assert(sizeof(Triangle) == 68, "reference Triangle layout");
_Static_assert(sizeof(Sphere) == 36, "reference Sphere layout");
_Static_assert(sizeof(Scene) == sizeof(OracleScene), "Scene layout");

/* the reference loaders are chatty on stdout; keep test logs readable */
static int hush_begin(void)
{
    fflush(stdout);
    int saved = dup(1);
    int devnull = open("/dev/null", O_WRONLY);
    if (devnull >= 0) {
        dup2(devnull, 1);
        close(devnull);
    }
    return saved;
}
static void hush_end(int saved)
{
    fflush(stdout);
    if (saved >= 0) {
        dup2(saved, 1);
        close(saved);
    }
}

/* loadOBJTriangles (raytracing.c:100) -> file-static triangles[]. The
   reference exit()s on a missing file, so existence is checked first. */
int ref_load_obj(const char *path)
{
    if (access(path, R_OK) != 0)
        return -1;
    int saved = hush_begin();
    loadOBJTriangles(path);
    hush_end(saved);
    return triangleCount;
}

/* parseTriangleFile (raytracing.c:76); writes <path>.parsed beside the input,
   so callers pass a writable copy. */
int ref_load_triangles_txt(const char *path)
{
    if (access(path, R_OK) != 0)
        return -1;
    triangleCount = 0;
    int saved = hush_begin();
    parseTriangleFile(path);
    hush_end(saved);
    return triangleCount;
}

/* hand an in-memory Triangle[] (68 B each) to the reference's closest-hit loop */
int ref_set_triangles(const void *tris, int n)
{
    triangles = malloc(sizeof(Triangle) * (size_t)(n > 0 ? n : 1)); /* the reference never frees either */
    memcpy(triangles, tris, sizeof(Triangle) * (size_t)n);
    triangleCount = n;
    return n;
}

int ref_triangle_count(void) { return triangleCount; }

int ref_get_triangles(void *out, int max)
{
    int n = triangleCount < max ? triangleCount : max;
    memcpy(out, triangles, sizeof(Triangle) * (size_t)n);
    return n;
}

int ref_sphere_count(void) { return sphereCount; }

int ref_get_spheres(void *out, int max)
{
    int n = sphereCount < max ? sphereCount : max;
    memcpy(out, spheres, sizeof(Sphere) * (size_t)n);
    return n;
}

/* camera basis with the reference's own normalized/cross, main.c:252-255 */
void ref_camera_look_at(const float origin[3], const float lookingAt[3], float fov, OracleCamera *out)
{
    vec3 o = {origin[0], origin[1], origin[2]};
    vec3 t = {lookingAt[0], lookingAt[1], lookingAt[2]};
    vec3 ez = normalized(minus(t, o));
    vec3 up = {0, -1, 0};
    vec3 ex = normalized(cross(ez, up));
    vec3 ey = normalized(cross(ez, ex));
    memcpy(&out->origin, &o, 12);
    memcpy(&out->ex, &ex, 12);
    memcpy(&out->ey, &ey, 12);
    memcpy(&out->ez, &ez, 12);
    out->fov = fov;
}

/* normalized() of the reference, for the sun direction (main.c:247) */
void ref_normalize(const float in[3], float out[3])
{
    vec3 v = {in[0], in[1], in[2]};
    v = normalized(v);
    out[0] = v.x;
    out[1] = v.y;
    out[2] = v.z;
}

/* index of the closest primary hit with the rule of calculateRayCollision
   (raytracing.c:216-237): spheres first, strict <, start 999999. */
static int32_t primary_id(Ray ray, int trianglesOnly)
{
    float best = 999999;
    int32_t id = -1;
    if (trianglesOnly == 0)
        for (int i = 0; i < sphereCount; ++i) {
            HitInfo h = raySphere(ray, spheres[i].pos, spheres[i].r);
            if (h.didHit && h.dst < best) {
                best = h.dst;
                id = -2 - i;
            }
        }
    for (int i = 0; i < triangleCount; ++i) {
        HitInfo h = rayTriangle(ray, triangles[i]);
        if (h.didHit && h.dst < best) {
            best = h.dst;
            id = i;
        }
    }
    return id;
}

/*
 * The pixel loop of main.c:84-101 with spp, size and pixel subset as
 * parameters. rgb: width*height*3 bytes, ids: width*height int32 (either may
 * be NULL); only the pixels of the subset are written.
 */
void ref_render(const OracleJob *job, uint8_t *rgb, int32_t *ids, OracleStats *stats)
{
    const int width = job->width, height = job->height;
    Scene scene;
    memcpy(&scene, &job->scene, sizeof(Scene));
    vec3 origin, ex, ey, ez;
    memcpy(&origin, &job->cam.origin, 12);
    memcpy(&ex, &job->cam.ex, 12);
    memcpy(&ey, &job->cam.ey, 12);
    memcpy(&ez, &job->cam.ez, 12);
    const float fov = job->cam.fov;
    const int spp = job->spp;
    const int G = job->nsplit > 1 ? job->nsplit : 1;
    const int rstep = job->row_step > 0 ? job->row_step : 1;
    const int cstep = job->col_step > 0 ? job->col_step : 1;
    uint64_t npix = 0, nhit = 0;

    for (int y = job->row_start; y < height; y += rstep) {
        for (int x = job->col_start; x < width; x += cstep) {
            float dx = (x - width / 2) / (float)(height / 2);
            float dy = (y - (height / 2)) / (float)(height / 2);
            vec3 dir = plus(plus(times(ex, dx), times(ey, dy)), times(ez, fov));
            dir = normalized(dir);
            Ray ray = {origin, dir};
            if (ids) {
                int32_t id = primary_id(ray, job->trianglesOnly);
                ids[x + y * width] = id;
                nhit += id != -1;
            }
            vec3 total = {0, 0, 0};
            for (int g = 0; g < G; ++g) {
                rngState = (unsigned)(x + y * width) + (unsigned)g * (unsigned)(width * height);
                vec3 acc = {0, 0, 0};
                int s0 = (int)((long long)g * spp / G), s1 = (int)((long long)(g + 1) * spp / G);
                for (int i = s0; i < s1; ++i)
                    acc = plus(acc, times(calcColor(ray, job->trianglesOnly, job->maxBounce, scene), 1. / spp));
                total = G == 1 ? acc : plus(total, acc);
            }
            if (rgb) {
                Color c = vec3ToColor(total);
                rgb[3 * (x + y * width) + 0] = c.r;
                rgb[3 * (x + y * width) + 1] = c.g;
                rgb[3 * (x + y * width) + 2] = c.b;
            }
            ++npix;
        }
    }
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->pixels = npix;
        stats->hit_pixels = nhit;
    }
}

/* single probes of reference functions, for unit tests of the restatement */
float ref_random_value(uint32_t *state)
{
    rngState = *state;
    float v = RandomValue();
    *state = rngState;
    return v;
}

void ref_random_direction(uint32_t *state, float out[3])
{
    rngState = *state;
    vec3 v = RandomDiretion();
    *state = rngState;
    out[0] = v.x;
    out[1] = v.y;
    out[2] = v.z;
}

/* the same for n independent states (test probe at scale) */
void ref_random_direction_batch(const uint32_t *states, int n, float *out, uint32_t *out_states)
{
    for (int i = 0; i < n; ++i) {
        uint32_t st = states[i];
        ref_random_direction(&st, out + 3 * (size_t)i);
        out_states[i] = st;
    }
}

void ref_environment_light(const float pos[3], const float dir[3], const OracleScene *s, float out[3])
{
    Ray r = {{pos[0], pos[1], pos[2]}, {dir[0], dir[1], dir[2]}};
    Scene sc;
    memcpy(&sc, s, sizeof sc);
    vec3 v = getEnvironmentLight(r, sc);
    out[0] = v.x;
    out[1] = v.y;
    out[2] = v.z;
}

/* n directions at once (test probe at scale) */
void ref_environment_light_batch(const float *dirs, int n, const OracleScene *s, float *out)
{
    const float pos[3] = {0.f, 0.f, 0.f};
    for (int i = 0; i < n; ++i)
        ref_environment_light(pos, dirs + 3 * (size_t)i, s, out + 3 * (size_t)i);
}

/* didHit/dst of one ray against triangle i / sphere k */
int ref_ray_triangle(const float pos[3], const float dir[3], int i, float *dst)
{
    Ray r = {{pos[0], pos[1], pos[2]}, {dir[0], dir[1], dir[2]}};
    HitInfo h = rayTriangle(r, triangles[i]);
    *dst = h.dst;
    return h.didHit;
}

int ref_ray_sphere(const float pos[3], const float dir[3], int k, float *dst, float normal[3])
{
    Ray r = {{pos[0], pos[1], pos[2]}, {dir[0], dir[1], dir[2]}};
    HitInfo h = raySphere(r, spheres[k].pos, spheres[k].r);
    *dst = h.dst;
    normal[0] = h.normal.x;
    normal[1] = h.normal.y;
    normal[2] = h.normal.z;
    return h.didHit;
}

uint8_t ref_float_to_uint(float f) { return floatToUint(f); }
