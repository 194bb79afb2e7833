/*
 * rayt_b200 -- host program of the B200 render loop: the reference's command line (reference main.c:47-79,
 * 119-231: same flag names, arity, defaults, "unknown argument -> usage, exit 0"), its scene loading modes
 * ("default" = triangles.txt + the sphere, or an .obj; main.c:233-244), its camera set-up (main.c:247-255) and its
 * image output (main.c:305), around ONE call of render_cuda() where the reference fans out 12 pthreads
 * (main.c:285-302).  New flags: --spp (the reference compiles 4000 in, scene.h:26), --gpus, --split, --nsplit.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "rtc_b200.h"
#include "rtc_host.h"

typedef struct {
    const char *s, *l; /* short and long spelling */
    int nargs;
    const char *what;
} Flag;

enum { F_HELP, F_IN, F_OUT, F_POS, F_TRACK, F_FOV, F_SIZE, F_BOUNCE, F_GC, F_SCH, F_SCZ, F_SUN, F_SPP, F_GPUS, F_SPLIT,
       F_NSPLIT, F_TRIS_TXT, F_COUNT };
static const Flag FLAGS[F_COUNT] = {
    {"-h", "--help", 0, ""},
    {"-i", "--input", 1, "string path/to/filename.obj"},
    {"-o", "--output", 1, "string outImage.bmp"},
    {"-p", "--pos", 3, "float posX posY posZ"},
    {"-t", "--track", 3, "float trackX trackY trackZ"},
    {"-f", "--fov", 1, "float fov"},
    {"-s", "--size", 2, "int width height"},
    {"-b", "--max-bounce", 1, "int maxBounce"},
    {"-gc", "--ground-color", 3, "float R G B"},
    {"-sch", "--sky-color-horizon", 3, "float R G B"},
    {"-scz", "--sky-color-zenith", 3, "float R G B"},
    {"--sun", "--sun", 5, "float x y z focus intensity"},
    {"--spp", "--spp", 1, "int samples per pixel"},
    {"--gpus", "--gpus", 1, "int number of GPUs"},
    {"--split", "--split", 1, "tiles|samples"},
    {"--nsplit", "--nsplit", 1, "int RNG sub-streams per pixel"},
    {"--triangles", "--triangles", 1, "path of the default scene file (triangles.txt)"},
};

static void usage(const char *prog, const char *err)
{
    if (err && *err)
        fprintf(stderr, "%s", err);
    printf("%s [-h|--help]\n"
           "   # FILE SETTINGS\n\t[-i|--input path/to/file.obj]\n\t[-o|--output <filename>]  (.png -> PNG, else BMP)\n"
           "   # POSITION & LOOKING-AT POSITION (TRACK)\n\t[-p|--pos <posX> <posY> <posZ>]\n\t[-t|--track <trackX> <trackY> <trackZ>]\n"
           "   # CAMERA SETTINGS\n\t[-f|--fov <fov>]\n\t[-s|--size <width> <height>]\n\t[-b|--max-bounce <maxBounce>]\n"
           "   # SCENE SETTINGS\n\t[-gc|--ground-color <R> <G> <B>]\n\t[-sch|--sky-color-horizon <R> <G> <B>]\n"
           "\t[-scz|--sky-color-zenith <R> <G> <B>]\n"
           "   # SUN SETTINGS\n\t[--sun <x> <y> <z> <focus> <intensity>]\n"
           "   # RENDER SETTINGS (new)\n\t[--spp <samples per pixel>]\n\t[--gpus <n>]\n"
           "\t[--split tiles|samples]  (tiles: the reference's image on any GPU count;\n"
           "\t                          samples: needs --nsplit, a multiple of --gpus -- its own RNG sub-streams)\n"
           "\t[--nsplit <sub-streams>]\n\t[--triangles path/to/triangles.txt]\n",
           prog);
    exit(0); /* the reference exits 0 from its usage text, errors included (main.c:63) */
}

int main(int argc, char **argv)
{
    /* defaults: reference main.c:10-28,111-116, scene.h:26 */
    const char *input = "default", *output = "out.bmp", *tri_file = "triangles.txt";
    float origin[3] = {-4.75f, -1.5f, -4.75f}, look[3] = {0.9f, -1.2f, 1.f}, fov = 1.f;
    int width = 128, height = 128, max_bounce = 10, spp = 4000, gpus = 1, split = RTC_SPLIT_TILES, nsplit = 1;
    float sun[3] = {-30.f, -85.f, 100.f};
    Scene scene = {{0, 0, 0}, {1.f, 1.f, 1.f}, {0.263f, 0.969f, 0.871f}, {.66f, .66f, .66f}, 22.f, .75f};

    for (int i = 1; i < argc; ++i) {
        int f = -1;
        for (int k = 0; k < F_COUNT; ++k)
            if (strcmp(argv[i], FLAGS[k].s) == 0 || strcmp(argv[i], FLAGS[k].l) == 0)
                f = k;
        if (f < 0) {
            char msg[512];
            snprintf(msg, sizeof msg, "ERROR: unknown argument '%s'\n", argv[i]);
            usage(argv[0], msg);
        }
        if (argc - 1 < i + FLAGS[f].nargs) {
            char msg[512];
            snprintf(msg, sizeof msg, "ERROR: %s/%s takes %d more param%s (%s)\n", FLAGS[f].l, FLAGS[f].s,
                     FLAGS[f].nargs, FLAGS[f].nargs > 1 ? "s" : "", FLAGS[f].what);
            usage(argv[0], msg);
        }
        char **v = argv + i + 1;
        switch (f) {
        case F_HELP: usage(argv[0], ""); break;
        case F_IN: input = v[0]; break;
        case F_OUT: output = v[0]; break;
        case F_POS: for (int k = 0; k < 3; ++k) origin[k] = (float)atof(v[k]); break;
        case F_TRACK: for (int k = 0; k < 3; ++k) look[k] = (float)atof(v[k]); break;
        case F_FOV: fov = (float)atof(v[0]); break;
        case F_SIZE: width = atoi(v[0]); height = atoi(v[1]); break;
        case F_BOUNCE: max_bounce = atoi(v[0]); break;
        case F_GC: scene.groundColor = (vec3){(float)atof(v[0]), (float)atof(v[1]), (float)atof(v[2])}; break;
        case F_SCH: scene.skyColorHorizon = (vec3){(float)atof(v[0]), (float)atof(v[1]), (float)atof(v[2])}; break;
        case F_SCZ: scene.skyColorZenith = (vec3){(float)atof(v[0]), (float)atof(v[1]), (float)atof(v[2])}; break;
        case F_SUN:
            for (int k = 0; k < 3; ++k) sun[k] = (float)atof(v[k]);
            scene.sunFocus = (float)atof(v[3]);
            scene.sunIntensity = (float)atof(v[4]);
            break;
        case F_SPP: spp = atoi(v[0]); break;
        case F_GPUS: gpus = atoi(v[0]); break;
        case F_SPLIT:
            if (strcmp(v[0], "tiles") == 0) split = RTC_SPLIT_TILES;
            else if (strcmp(v[0], "samples") == 0) split = RTC_SPLIT_SAMPLES;
            else usage(argv[0], "ERROR: --split takes tiles or samples\n");
            break;
        case F_NSPLIT: nsplit = atoi(v[0]); break;
        case F_TRIS_TXT: tri_file = v[0]; break;
        }
        i += FLAGS[f].nargs;
    }

    /* scene: "default" = triangles.txt + the hard-coded sphere, else an OBJ with trianglesOnly (main.c:233-244) */
    Triangle *tris = NULL;
    int ntris = 0, triangles_only = 0, rc;
    printf("Mode=%s\n", input);
    if (strcmp(input, "default") == 0) {
        rc = host_load_triangles_txt(tri_file, &tris, &ntris);
    } else {
        triangles_only = 1;
        rc = host_load_obj(input, &tris, &ntris);
    }
    if (rc != HOST_OK) {
        fprintf(stderr, "Error while loading '%s' (code %d)\n", strcmp(input, "default") ? input : tri_file, rc);
        return rc == HOST_ERR_FORMAT ? 69 : 42; /* the reference's exit codes */
    }
    printf("%d triangles found\n", ntris);
    int nspheres = 0;
    const Sphere *spheres = host_default_spheres(&nspheres);

    /* camera and sun: main.c:247-255 */
    float nsun[3];
    rtc_normalize(sun, nsun);
    scene.normalizedSunDirection = (vec3){nsun[0], nsun[1], nsun[2]};
    Camera cam;
    rtc_camera_look_at(origin, look, fov, &cam);

    uint8_t *image = malloc((size_t)width * height * 3); /* heap: the reference's stack VLA overflows at 4K */
    if (!image)
        return 1;
    RenderStats st;
    RenderParams rp;
    memset(&rp, 0, sizeof rp);
    rp.tris = tris; rp.ntris = ntris;
    rp.spheres = spheres; rp.nspheres = nspheres;
    rp.width = width; rp.height = height; rp.spp = spp; rp.maxBounce = max_bounce;
    rp.trianglesOnly = triangles_only;
    rp.ngpus = gpus; rp.split_mode = split; rp.nsplit = nsplit;
    rp.stats = &st;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    rc = render_cuda(&scene, &cam, image, &rp); /* replaces the pthread fan-out of main.c:285-302 */
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (rc != RTC_OK) {
        fprintf(stderr, "render_cuda failed (%d): %s\n", rc, rtc_last_error());
        return 1;
    }
    double wall = (t1.tv_sec - t0.tv_sec) * 1e3 + (t1.tv_nsec - t0.tv_nsec) * 1e-6;
    printf("%dx%d, %d spp, %d bounces on %d GPU(s): %llu rays, kernels %.3f ms, call %.3f ms (%.1f Mrays/s)\n", width,
           height, spp, max_bounce, st.ngpus_used, (unsigned long long)st.rays, st.kernel_ms, wall,
           st.rays / (wall * 1e-3) / 1e6);
    if (host_write_image(output, width, height, image) != HOST_OK) {
        fprintf(stderr, "cannot write %s\n", output);
        return 1;
    }
    printf("wrote %s\n", output);
    render_cuda_shutdown();
    free(image);
    free(tris);
    return 0;
}
