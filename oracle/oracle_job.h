/*
 * TEST INFRASTRUCTURE -- not product code.
 *
 * Job/stat structs shared by the two CPU checkers in this directory:
 *   ref_harness.c  : the reference's own C functions, compiled from
 *                    /root/reference where they lie, into oracle/_ref/
 *   rtc_oracle.c   : an independent plain-C restatement of the same algorithm
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
 * load anything from oracle/.
 */
#ifndef ORACLE_JOB_H
#define ORACLE_JOB_H

#include <stdint.h>

typedef struct OracleVec3 { float x, y, z; } OracleVec3;

typedef struct OracleScene { /* == reference Scene, raytracing.h:7-11 */
    OracleVec3 sun, horizon, zenith, ground;
    float sunFocus, sunIntensity;
} OracleScene;

typedef struct OracleCamera { /* == Camera in include/rtc_types.h */
    OracleVec3 origin, ex, ey, ez;
    float fov;
} OracleCamera;

typedef struct OracleJob {
    int32_t width, height;
    int32_t spp;           /* reference: compile-time accumulationCount, scene.h:26 */
    int32_t maxBounce;     /* reference: -b flag, main.c:168-174 */
    int32_t trianglesOnly; /* reference: main.c:113,241 */
    /* pixel subset: rows row_start, row_start+row_step, ...; same for columns.
       Every pixel is independent (own seed), so any subset is exact. */
    int32_t row_start, row_step;
    int32_t col_start, col_step;
    /* nsplit = 1 is the reference. nsplit = G > 1 is the sample-split
       definition of DESIGN.md: sub-stream g is seeded x + y*W + g*W*H, renders
       samples [g*spp/G, (g+1)*spp/G) into its own partial sum, partial sums
       are added in g order, then quantised. */
    int32_t nsplit;
    OracleScene scene;
    OracleCamera cam;
} OracleJob;

typedef struct OracleStats {
    uint64_t pixels;        /* pixels rendered */
    uint64_t hit_pixels;    /* pixels whose primary ray hit something */
    uint64_t rays;          /* calls of the closest-hit loop (raytracing.c:270) */
    uint64_t sphere_tests;  /* ray-sphere pairs */
    /* ray-triangle pairs by the reference's own exit (raytracing.c:189-213):
       [0] cull  [1] det  [2] u  [3] v / u+v  [4] dst<EPS  [5] hit */
    uint64_t tri_exit[6];
    uint64_t rng_draws;     /* RandomValue() calls */
    uint64_t sky_evals;     /* getEnvironmentLight() calls */
} OracleStats;

#endif
