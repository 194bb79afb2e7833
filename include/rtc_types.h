/*
 * rtc_types.h -- plain-C data model shared by the host program, the C-ABI
 * render library (rtc_b200.h) and the test oracle.
 *
 * The layouts are binary-compatible with the reference's structs so that a
 * reference build can hand its arrays to render_cuda() unchanged:
 *
 *   vec3      reference moremath.h:10-13      12 B  {x,y,z}
 *   Scene     reference raytracing.h:7-11     56 B  sky / sun parameters only
 *   Color     reference raytracing.h:15-18     3 B  {r,g,b}
 *   Material  reference raytracing.h:25-30    20 B
 *   Sphere    reference raytracing.h:34-39    36 B
 *   Triangle  reference raytracing.h:41-45    68 B
 *
 * When this header is included next to the reference's own headers define
 * RTC_TYPES_FROM_REFERENCE first: the names above are then taken from the
 * reference and only the new types (Camera, ...) are declared here.
 */
#ifndef RTC_TYPES_H
#define RTC_TYPES_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef RTC_TYPES_FROM_REFERENCE

typedef struct vec3 {
    float x, y, z;
} vec3;

typedef struct Scene {
    vec3 normalizedSunDirection;
    vec3 skyColorHorizon;
    vec3 skyColorZenith;
    vec3 groundColor;
    float sunFocus;
    float sunIntensity;
} Scene;

typedef struct Color {
    uint8_t r, g, b;
} Color;

typedef struct Material {
    vec3 color;
    float emissionStrength;
    float smoothness;
} Material;

typedef struct Sphere {
    vec3 pos;
    float r;
    Material mat;
} Sphere;

typedef struct Triangle {
    vec3 posA, posB, posC;
    vec3 normal; /* as stored: NOT recomputed, NOT renormalised by the renderer */
    Material mat;
} Triangle;

#endif /* RTC_TYPES_FROM_REFERENCE */

/*
 * Camera: the five values the reference passes to each render thread
 * (threadArgs{origin, ex, ey, ez, fov}, reference main.c:30-38); the basis is
 * built on the host exactly as main.c:252-255 does (rtc_camera_look_at).
 * fov is a multiplier on ez (a focal length), not an angle (main.c:90).
 */
typedef struct Camera {
    vec3 origin;
    vec3 ex, ey, ez;
    float fov;
} Camera;

/* id written to the primary-hit map */
#define RTC_ID_MISS (-1)
#define RTC_ID_SPHERE(k) (-2 - (k))

#ifdef __cplusplus
}
#endif

#endif /* RTC_TYPES_H */
