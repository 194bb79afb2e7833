/*
 * rtc_host.h -- host-side C of the renderer: scene ingestion and image output.
 *
 * Written fresh for this repository; it reproduces the OBSERVABLE behaviour of the reference's loaders so the
 * Triangle[] handed to render_cuda() is the array the reference would have built:
 *   host_load_obj            loadObj (reference objloader.c:340-551) + loadOBJTriangles (raytracing.c:100-147)
 *   host_load_triangles_txt  parseTriangleFile / cleanFile (raytracing.c:19-98), without the ".parsed" side file
 *   host_write_bmp           what stbi_write_bmp(name, w, h, 3, data) produces (reference main.c:305)
 *   host_write_png           8-bit RGB PNG (BASELINE.json north_star asks for PNG output)
 */
#ifndef RTC_HOST_H
#define RTC_HOST_H

#include <stdint.h>

#include "rtc_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define HOST_OK 0
#define HOST_ERR_OPEN 1        /* file not found / unreadable (reference: exit(42), raytracing.c:109)   */
#define HOST_ERR_FORMAT 2      /* face without texture index "v//vn" (reference: exit(69), objloader.c:521) */
#define HOST_ERR_NOMEM 3

/* Wavefront OBJ (+ MTL next to it) -> malloc'ed Triangle[]; *ntris faces in file order.  Caller frees. */
int host_load_obj(const char *path, Triangle **tris, int *ntris);
/* the reference's "default" scene format (triangles.txt) -> malloc'ed Triangle[] */
int host_load_triangles_txt(const char *path, Triangle **tris, int *ntris);
/* the one sphere the reference hard-codes (scene.h:17-19) */
const Sphere *host_default_spheres(int *count);

/* rgb: width*height*3 bytes, row 0 on top (the reference's Color image[]) */
int host_write_bmp(const char *path, int width, int height, const uint8_t *rgb);
int host_write_png(const char *path, int width, int height, const uint8_t *rgb);
/* picks by extension: ".png" -> PNG, anything else -> BMP (the reference always writes BMP) */
int host_write_image(const char *path, int width, int height, const uint8_t *rgb);

#ifdef __cplusplus
}
#endif
#endif
