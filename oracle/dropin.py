"""dropin.py -- proves the drop-in claim of INTEGRATION.md against the reference's OWN main.c.

TEST INFRASTRUCTURE (like everything under oracle/): copies the reference sources from /root/reference to a scratch
directory, applies the patch INTEGRATION.md section 1 describes (an accessor for raytracing.c's private triangle
array; main.c:246 -> heap image; main.c:285-302, the pthread fan-out, -> one render_cuda() call), and builds the
result with gcc against include/rtc_b200.h and raytracingc_b200/librtc_b200.so.  Nothing of the reference is copied
into the repository: the binary and the one model file the GPU test feeds it go to oracle/_ref/dropin/ (git-ignored,
travels to the GPU box like the other built files).

    python -m oracle.dropin            # build; prints the binary's path
"""
from __future__ import annotations

import os
import shutil
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
OUT_DIR = os.path.join(HERE, "_ref", "dropin")
BINARY = os.path.join(OUT_DIR, "rayt_dropin")
STOCK = os.path.join(OUT_DIR, "rayt_stock")  # the UNMODIFIED reference program (12 pthreads), timed for the record

# what replaces main.c:285-302
CALL = r'''
   {
      Camera rtc_cam = {origin, ex, ey, ez, fov};
      int rtc_ntris = 0;
      const Triangle *rtc_tris = getTriangles(&rtc_ntris);
      RenderParams rp;
      memset(&rp, 0, sizeof rp);
      rp.tris = rtc_tris;
      rp.ntris = rtc_ntris;
      rp.spheres = spheres;
      rp.nspheres = sphereCount;
      rp.width = width;
      rp.height = height;
      rp.spp = accumulationCount;
      rp.maxBounce = maxBounce;
      rp.trianglesOnly = trianglesOnly;
      rp.ngpus = 1;
      rp.split_mode = RTC_SPLIT_TILES;
      if (render_cuda(&scene, &rtc_cam, (uint8_t *)image, &rp) != RTC_OK)
      {
         fprintf(stderr, "render_cuda: %s\n", rtc_last_error());
         return 1;
      }
   }
'''


def _replace_once(text: str, old: str, new: str, what: str) -> str:
    assert text.count(old) == 1, f"dropin: anchor for {what} found {text.count(old)} times in the reference"
    return text.replace(old, new)


def patch_sources(src_dir: str) -> None:
    """Apply the INTEGRATION.md patch to a scratch copy of the reference."""
    # accessor next to loadOBJTriangles (raytracing.c keeps the only filled copy of the header-static array)
    p = os.path.join(src_dir, "raytracing.c")
    with open(p) as f:
        t = f.read()
    t += "\nconst Triangle *getTriangles(int *count) { *count = triangleCount; return triangles; }\n"
    with open(p, "w") as f:
        f.write(t)
    p = os.path.join(src_dir, "raytracing.h")
    with open(p) as f:
        t = f.read()
    t = _replace_once(t, "void loadOBJTriangles(char const *filename);",
                      "void loadOBJTriangles(char const *filename);\nconst Triangle *getTriangles(int *count);", "raytracing.h")
    with open(p, "w") as f:
        f.write(t)
    p = os.path.join(src_dir, "main.c")
    with open(p) as f:
        t = f.read()
    t = _replace_once(t, '#include "scene.h"\n',
                      '#include "scene.h"\n#define RTC_TYPES_FROM_REFERENCE\n#include "rtc_b200.h"\n', "the include")
    t = _replace_once(t, "   Color image[width * height];\n",
                      "   Color *image = malloc((size_t)width * height * sizeof(Color));\n", "main.c:246")
    a = t.index("   pthread_t tIDs[NUMBER_OF_THREADS];")
    b = t.index("#endif\n   stbi_write_bmp(")
    t = t[:a] + CALL + t[b:]  # main.c:285-302 -> render_cuda()
    with open(p, "w") as f:
        f.write(t)


def build(verbose: bool = False) -> str:
    """Patch a scratch copy of the reference and link it against librtc_b200.so; returns the binary's path."""
    if not os.path.exists(os.path.join(REF, "main.c")):
        raise FileNotFoundError(f"{REF} is not available (it only exists in the build container)")
    lib_dir = os.path.join(ROOT, "raytracingc_b200")
    if not os.path.exists(os.path.join(lib_dir, "librtc_b200.so")):
        raise FileNotFoundError("librtc_b200.so not built: run python -m raytracingc_b200.build")
    os.makedirs(os.path.join(OUT_DIR, "3Dmodels"), exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="rtc_dropin_") as tmp:
        src = os.path.join(tmp, "ref")
        shutil.copytree(REF, src, ignore=shutil.ignore_patterns(".git", "*.o", "*.out", "*.bmp"))
        # the stock program first, from the untouched copy (the reference's own Makefile puts -lm before the objects and
        # fails to link on this toolchain, SURVEY fact 9: same sources, same -O3, libraries at the end)
        subprocess.run(["gcc", "-O3", "-w", "main.c", "raytracing.c", "moremath.c", "objloader.c", "-o", STOCK, "-lm", "-lpthread"],
                       cwd=src, check=True)
        patch_sources(src)
        cmd = ["gcc", "-O3", "-w", "-I" + os.path.join(ROOT, "include"), "main.c", "raytracing.c", "moremath.c", "objloader.c",
               "-o", BINARY, "-L" + lib_dir, "-lrtc_b200", "-Wl,-rpath,$ORIGIN/../../../raytracingc_b200", "-lm", "-lpthread"]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, cwd=src, check=True)
        for name in ("simplest.obj", "simplest.mtl"):  # the model of the one vector the stock binary reproduces too
            if os.path.exists(os.path.join(src, "3Dmodels", name)):
                dst = os.path.join(OUT_DIR, "3Dmodels", name)
                if os.path.exists(dst):
                    os.chmod(dst, 0o644)
                shutil.copyfile(os.path.join(src, "3Dmodels", name), dst)
    return BINARY


if __name__ == "__main__":
    print(build(verbose=True))
