/* image_out.c -- BMP (byte-compatible with what the reference's writer call produces) and PNG output. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rtc_host.h"

static void put16(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); }
static void put32(uint8_t *p, uint32_t v) { put16(p, v); put16(p + 2, v >> 16); }
static void put32be(uint8_t *p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = (uint8_t)v; }

/* 24-bit uncompressed BMP: 14-byte file header + 40-byte info header, rows bottom-up, BGR, padded to 4 bytes;
 * the size / resolution / palette fields are zero -- the file stbi_write_bmp(name, w, h, 3, data) writes
 * (reference main.c:305). */
int host_write_bmp(const char *path, int w, int h, const uint8_t *rgb)
{
    if (!path || !rgb || w <= 0 || h <= 0)
        return HOST_ERR_FORMAT;
    FILE *fp = fopen(path, "wb");
    if (!fp)
        return HOST_ERR_OPEN;
    const int pad = (-w * 3) & 3, stride = w * 3 + pad;
    uint8_t hdr[54];
    memset(hdr, 0, sizeof hdr);
    hdr[0] = 'B';
    hdr[1] = 'M';
    put32(hdr + 2, (uint32_t)(54 + (size_t)stride * h));
    put32(hdr + 10, 54);
    put32(hdr + 14, 40);
    put32(hdr + 18, (uint32_t)w);
    put32(hdr + 22, (uint32_t)h);
    put16(hdr + 26, 1);
    put16(hdr + 28, 24);
    fwrite(hdr, 1, sizeof hdr, fp);
    uint8_t *row = calloc((size_t)stride, 1);
    if (!row) {
        fclose(fp);
        return HOST_ERR_NOMEM;
    }
    for (int y = h - 1; y >= 0; --y) {
        const uint8_t *src = rgb + (size_t)y * w * 3;
        for (int x = 0; x < w; ++x) {
            row[3 * x] = src[3 * x + 2];
            row[3 * x + 1] = src[3 * x + 1];
            row[3 * x + 2] = src[3 * x];
        }
        fwrite(row, 1, (size_t)stride, fp);
    }
    free(row);
    return fclose(fp) == 0 ? HOST_OK : HOST_ERR_OPEN;
}

/* ---- PNG: 8-bit RGB, filter 0 rows, zlib stream of stored (uncompressed) deflate blocks ---- */
static uint32_t crc_table[256];
static void crc_init(void)
{
    for (uint32_t n = 0; n < 256; ++n) {
        uint32_t c = n;
        for (int k = 0; k < 8; ++k)
            c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
}
static uint32_t crc_update(uint32_t c, const uint8_t *p, size_t n)
{
    for (size_t i = 0; i < n; ++i)
        c = crc_table[(c ^ p[i]) & 0xff] ^ (c >> 8);
    return c;
}
static void chunk(FILE *fp, const char type[4], const uint8_t *data, uint32_t len)
{
    uint8_t b[4];
    put32be(b, len);
    fwrite(b, 1, 4, fp);
    fwrite(type, 1, 4, fp);
    if (len)
        fwrite(data, 1, len, fp);
    uint32_t c = crc_update(0xffffffffu, (const uint8_t *)type, 4);
    c = crc_update(c, data, len) ^ 0xffffffffu;
    put32be(b, c);
    fwrite(b, 1, 4, fp);
}

int host_write_png(const char *path, int w, int h, const uint8_t *rgb)
{
    if (!path || !rgb || w <= 0 || h <= 0)
        return HOST_ERR_FORMAT;
    if (!crc_table[1])
        crc_init();
    const size_t row = (size_t)w * 3 + 1, raw = row * (size_t)h;
    const size_t nblocks = (raw + 65534) / 65535;
    const size_t zlen = 2 + raw + 5 * nblocks + 4;
    if (zlen > 0x7fffffffu)
        return HOST_ERR_FORMAT; /* one IDAT chunk holds the whole image */
    uint8_t *z = malloc(zlen), *scan = malloc(raw);
    if (!z || !scan) {
        free(z); free(scan);
        return HOST_ERR_NOMEM;
    }
    uint32_t s1 = 1, s2 = 0; /* adler32 of the raw scanlines */
    for (int y = 0; y < h; ++y) {
        scan[(size_t)y * row] = 0; /* filter type 0 */
        memcpy(scan + (size_t)y * row + 1, rgb + (size_t)y * w * 3, (size_t)w * 3);
    }
    for (size_t i = 0; i < raw; ++i) {
        s1 = (s1 + scan[i]) % 65521u;
        s2 = (s2 + s1) % 65521u;
    }
    size_t o = 0;
    z[o++] = 0x78;
    z[o++] = 0x01;
    for (size_t b = 0, pos = 0; b < nblocks; ++b) {
        const size_t n = raw - pos < 65535 ? raw - pos : 65535;
        z[o++] = (b + 1 == nblocks) ? 1 : 0; /* BFINAL, BTYPE = 00 (stored) */
        put16(z + o, (uint32_t)n);
        put16(z + o + 2, (uint32_t)~n);
        o += 4;
        memcpy(z + o, scan + pos, n);
        o += n;
        pos += n;
    }
    put32be(z + o, (s2 << 16) | s1);
    o += 4;
    free(scan);
    FILE *fp = fopen(path, "wb");
    if (!fp) {
        free(z);
        return HOST_ERR_OPEN;
    }
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    fwrite(sig, 1, 8, fp);
    uint8_t ihdr[13];
    put32be(ihdr, (uint32_t)w);
    put32be(ihdr + 4, (uint32_t)h);
    ihdr[8] = 8;  /* bit depth */
    ihdr[9] = 2;  /* colour type: RGB */
    ihdr[10] = ihdr[11] = ihdr[12] = 0;
    chunk(fp, "IHDR", ihdr, 13);
    chunk(fp, "IDAT", z, (uint32_t)o);
    chunk(fp, "IEND", NULL, 0);
    free(z);
    return fclose(fp) == 0 ? HOST_OK : HOST_ERR_OPEN;
}

int host_write_image(const char *path, int w, int h, const uint8_t *rgb)
{
    const char *dot = path ? strrchr(path, '.') : NULL;
    if (dot && (strcmp(dot, ".png") == 0 || strcmp(dot, ".PNG") == 0))
        return host_write_png(path, w, h, rgb);
    return host_write_bmp(path, w, h, rgb);
}
