/* loader.c -- OBJ/MTL and triangles.txt ingestion with the reference loaders' observable behaviour (rtc_host.h). */
#define _GNU_SOURCE
#include <libgen.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rtc_host.h"

typedef struct {
    char name[128];
    float color[3];
    float emission, smoothness;
} Mtl;

/* one sphere: centre (0,1,0), radius 2.5, white, no emission, matte (reference scene.h:17-19) */
static const Sphere k_spheres[1] = {{{0.f, 1.f, 0.f}, 2.5f, {{1.f, 1.f, 1.f}, 0.f, 0.f}}};
const Sphere *host_default_spheres(int *count)
{
    if (count)
        *count = 1;
    return k_spheres;
}

static int starts(const char *line, const char *key) { return strncmp(line, key, strlen(key)) == 0; }

/* skipped lines: empty ("\n" only) and comments (objloader.c:397) */
static int skip_line(const char *line, size_t len) { return len == 1 || line[0] == '#'; }

/* MTL: "newmtl name", "Ns f" -> smoothness = sqrt(0.001 * Ns) in double (objloader.c:272), "Kd r g b" -> colour,
 * "Ke e _ _" -> emission = FIRST component (objloader.c:297).  A key a material lacks stays 0 here (the reference
 * leaves malloc memory uninitialised, objloader.c:150). */
static int load_mtl(const char *path, Mtl **out, int *count)
{
    FILE *fp = fopen(path, "r");
    *out = NULL;
    *count = 0;
    if (!fp)
        return HOST_ERR_OPEN; /* reference: warning, default material for every face (objloader.c:234-240) */
    char *line = NULL;
    size_t cap = 0;
    ssize_t len;
    int n = 0;
    while ((len = getline(&line, &cap, fp)) != -1)
        if (!skip_line(line, (size_t)len) && starts(line, "newmtl "))
            n++;
    Mtl *m = calloc((size_t)(n > 0 ? n : 1), sizeof(Mtl));
    if (!m) {
        fclose(fp);
        free(line);
        return HOST_ERR_NOMEM;
    }
    rewind(fp);
    int cur = -1;
    while ((len = getline(&line, &cap, fp)) != -1) {
        if (skip_line(line, (size_t)len))
            continue;
        if (starts(line, "newmtl ")) {
            char name[256];
            if (sscanf(line, "newmtl %255s", name) == 1 && cur + 1 < n) {
                cur++;
                size_t nl = strlen(name);
                if (nl > sizeof m[cur].name - 1)
                    nl = sizeof m[cur].name - 1; /* the fixed-size field truncates long names */
                memcpy(m[cur].name, name, nl);
                m[cur].name[nl] = '\0';
            }
        }
        if (cur < 0)
            continue; /* a key before any newmtl: the reference writes out of bounds; ignored here */
        float a, b, c;
        if (starts(line, "Ns ") && sscanf(line, "Ns %f", &a) == 1)
            m[cur].smoothness = (float)sqrt(0.001 * a);
        if (starts(line, "Kd ") && sscanf(line, "Kd %f %f %f", &a, &b, &c) >= 1) {
            /* sscanf fills as many components as it finds, like the reference's direct scan into the struct */
            int got = sscanf(line, "Kd %f %f %f", &m[cur].color[0], &m[cur].color[1], &m[cur].color[2]);
            (void)got;
        }
        if (starts(line, "Ke ") && sscanf(line, "Ke %f %f %f", &a, &b, &c) >= 1)
            m[cur].emission = a;
    }
    free(line);
    fclose(fp);
    *out = m;
    *count = n;
    return HOST_OK;
}

int host_load_obj(const char *path, Triangle **tris_out, int *ntris_out)
{
    *tris_out = NULL;
    *ntris_out = 0;
    FILE *fp = fopen(path, "r");
    if (!fp)
        return HOST_ERR_OPEN;
    char *line = NULL;
    size_t cap = 0;
    ssize_t len;
    /* pass 1: count (objloader.c:157-195) */
    int nv = 0, nn = 0, nf = 0;
    while ((len = getline(&line, &cap, fp)) != -1) {
        if (skip_line(line, (size_t)len))
            continue;
        if (starts(line, "v "))
            nv++;
        if (starts(line, "vn "))
            nn++;
        else if (starts(line, "f "))
            nf++;
    }
    float(*v)[3] = calloc((size_t)nv + 1, sizeof *v);   /* 1-based, objloader.c:205-207 */
    float(*vn)[3] = calloc((size_t)nn + 1, sizeof *vn);
    Triangle *tris = calloc((size_t)(nf > 0 ? nf : 1), sizeof(Triangle));
    if (!v || !vn || !tris) {
        free(v); free(vn); free(tris); free(line);
        fclose(fp);
        return HOST_ERR_NOMEM;
    }
    char dirbuf[1024];
    snprintf(dirbuf, sizeof dirbuf, "%s", path);
    const char *dir = dirname(dirbuf);
    char dircopy[1024];
    snprintf(dircopy, sizeof dircopy, "%s", dir);

    Mtl *mtls = NULL;
    int nmtl = 0, cur_mtl = -1; /* persists across objects until the next usemtl (objloader.c:51) */
    int iv = 0, in = 0, it = 0, rc = HOST_OK;
    rewind(fp);
    while ((len = getline(&line, &cap, fp)) != -1) {
        if (skip_line(line, (size_t)len))
            continue;
        if (starts(line, "mtllib ")) {
            char name[256];
            if (sscanf(line, "mtllib %255s", name) == 1) {
                char mpath[1400];
                snprintf(mpath, sizeof mpath, "%s/%s", dircopy, name);
                free(mtls);
                if (load_mtl(mpath, &mtls, &nmtl) != HOST_OK)
                    fprintf(stderr, "warning: material file %s not found, using the default material\n", mpath);
            }
        }
        if (starts(line, "usemtl ")) {
            char name[256];
            if (sscanf(line, "usemtl %255s", name) == 1) {
                cur_mtl = -1; /* unknown name -> default white (objloader.c:413-430) */
                for (int i = 0; i < nmtl; ++i)
                    if (strcmp(name, mtls[i].name) == 0) {
                        cur_mtl = i;
                        break;
                    }
            }
        }
        if (starts(line, "v ") && iv < nv) {
            iv++;
            sscanf(line, "v %f %f %f", &v[iv][0], &v[iv][1], &v[iv][2]);
        }
        if (starts(line, "vn ") && in < nn) {
            in++;
            sscanf(line, "vn %f %f %f", &vn[in][0], &vn[in][1], &vn[in][2]);
        }
        if (starts(line, "f ") && it < nf) {
            int av, at, an, bv, bt, bn, cv, ct, cn;
            /* only "a/b/c d/e/f g/h/i" is understood; further vertices of a polygon are ignored, so a quad
             * contributes its first triangle only (objloader.c:486) */
            int got = sscanf(line, "f %d/%d/%d %d/%d/%d %d/%d/%d", &av, &at, &an, &bv, &bt, &bn, &cv, &ct, &cn);
            if (got == 9 && av >= 0 && bv >= 0 && cv >= 0 && an >= 0 && av <= nv && bv <= nv && cv <= nv && an <= nn) {
                Triangle *t = &tris[it];
                /* rotate by 180 degrees about z: (x, y, z) -> (-x, -y, z), positions and the normal
                 * (raytracing.c:123-135); the normal is vertex A's vn, neither recomputed nor renormalised */
                t->posA = (vec3){-v[av][0], -v[av][1], v[av][2]};
                t->posB = (vec3){-v[bv][0], -v[bv][1], v[bv][2]};
                t->posC = (vec3){-v[cv][0], -v[cv][1], v[cv][2]};
                t->normal = (vec3){-vn[an][0], -vn[an][1], vn[an][2]};
                if (cur_mtl < 0 || cur_mtl >= nmtl) {
                    t->mat.color = (vec3){1.f, 1.f, 1.f};
                    t->mat.emissionStrength = 0.f;
                    t->mat.smoothness = 0.f;
                } else {
                    t->mat.color = (vec3){mtls[cur_mtl].color[0], mtls[cur_mtl].color[1], mtls[cur_mtl].color[2]};
                    t->mat.emissionStrength = mtls[cur_mtl].emission;
                    t->mat.smoothness = mtls[cur_mtl].smoothness;
                }
            } else if (got >= 1 && strstr(line, "//")) {
                rc = HOST_ERR_FORMAT; /* "v//vn" faces: the reference gives up (objloader.c:518-521) */
                break;
            }
            it++;
        }
    }
    free(line);
    free(v);
    free(vn);
    free(mtls);
    fclose(fp);
    if (rc != HOST_OK) {
        free(tris);
        return rc;
    }
    *tris_out = tris;
    *ntris_out = nf;
    return HOST_OK;
}

/* triangles.txt: everything except digits, '-', '.', '+', newline becomes a space, "//" comments run to the end
 * of the line, a lone '/' is dropped (cleanFile, raytracing.c:47-74); then a count and 14 floats per triangle
 * (A, B, C, colour, emission, smoothness); normal = normalized(cross(B-A, C-A)) in the reference's float
 * arithmetic (raytracing.c:24). */
int host_load_triangles_txt(const char *path, Triangle **tris_out, int *ntris_out)
{
    *tris_out = NULL;
    *ntris_out = 0;
    FILE *fp = fopen(path, "r");
    if (!fp)
        return HOST_ERR_OPEN;
    fseek(fp, 0, SEEK_END);
    long size = ftell(fp);
    rewind(fp);
    char *raw = malloc((size_t)size + 1), *clean = malloc((size_t)size + 2);
    if (!raw || !clean) {
        free(raw); free(clean);
        fclose(fp);
        return HOST_ERR_NOMEM;
    }
    size_t n = fread(raw, 1, (size_t)size, fp), o = 0;
    fclose(fp);
    for (size_t i = 0; i < n; ++i) {
        char c = raw[i];
        if ((c >= '0' && c <= '9') || c == '-' || c == '.' || c == '\n' || c == '+') {
            clean[o++] = c;
        } else if (c == '/') {
            if (i + 1 < n && raw[i + 1] == '/') {
                while (i < n && raw[i] != '\n')
                    ++i; /* the newline itself is consumed too, as skipUntilDelim does */
            }
        } else {
            clean[o++] = ' ';
        }
    }
    clean[o] = 0;
    free(raw);
    char *p = clean, *end;
    long count = strtol(p, &end, 0); /* "%i" */
    if (end == p || count < 0) {
        free(clean);
        return HOST_ERR_FORMAT;
    }
    p = end;
    Triangle *tris = calloc((size_t)(count > 0 ? count : 1), sizeof(Triangle));
    if (!tris) {
        free(clean);
        return HOST_ERR_NOMEM;
    }
    for (long i = 0; i < count; ++i) {
        float f[14] = {0};
        for (int k = 0; k < 14; ++k) {
            f[k] = strtof(p, &end);
            if (end == p)
                break; /* short file: remaining fields stay 0 (fscanf would leave them untouched) */
            p = end;
        }
        Triangle *t = &tris[i];
        t->posA = (vec3){f[0], f[1], f[2]};
        t->posB = (vec3){f[3], f[4], f[5]};
        t->posC = (vec3){f[6], f[7], f[8]};
        t->mat.color = (vec3){f[9], f[10], f[11]};
        t->mat.emissionStrength = f[12];
        t->mat.smoothness = f[13];
        /* minus, cross, normalized of reference moremath.c, one IEEE float operation each (-ffp-contract=off) */
        float ab[3] = {f[3] - f[0], f[4] - f[1], f[5] - f[2]}, ac[3] = {f[6] - f[0], f[7] - f[1], f[8] - f[2]};
        float nx = ab[1] * ac[2] - ab[2] * ac[1], ny = ab[2] * ac[0] - ab[0] * ac[2], nz = ab[0] * ac[1] - ab[1] * ac[0];
        float len = (float)sqrt((double)(nx * nx + ny * ny + nz * nz));
        float inv = (float)(1. / (double)len);
        t->normal = (vec3){nx * inv, ny * inv, nz * inv};
    }
    free(clean);
    *tris_out = tris;
    *ntris_out = (int)count;
    return HOST_OK;
}
