// Image output for the RGBA8 result of render() — SURVEY §8(f).4.  The reference never saves its image (src/cli.rs:44-46 drops
// it; the GUI only displays it), so there is no reference format to match: PNG (8-bit RGBA, zlib "stored" blocks — no compressor
// dependency) and binary PPM (alpha dropped).  Headless: bytes to a file, no window.
#include <cstdio>
#include <cstring>
#include <vector>

#include "host_common.h"

using namespace mp;

static uint32_t crc_table[256];
static bool crc_ready = false;
static uint32_t crc32_update(uint32_t c, const uint8_t* p, size_t n) {
    if (!crc_ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t v = i;
            for (int k = 0; k < 8; ++k) v = (v & 1u) ? 0xEDB88320u ^ (v >> 1) : v >> 1;
            crc_table[i] = v;
        }
        crc_ready = true;
    }
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
    return c;
}
static void put32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back(x >> 24), v.push_back(x >> 16), v.push_back(x >> 8), v.push_back(x);
}
static void chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
    put32(out, (uint32_t)data.size());
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put32(out, crc32_update(0xFFFFFFFFu, &out[at], 4 + data.size()) ^ 0xFFFFFFFFu);
}

extern "C" {

int mp_image_write_png(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height) {
    if (!path || !rgba || width == 0 || height == 0) return fail(MP_EINVAL, "bad argument");
    std::vector<uint8_t> raw;  // filter byte 0 + row
    raw.reserve((size_t)height * (1 + (size_t)width * 4));
    for (uint32_t y = 0; y < height; ++y) {
        raw.push_back(0);
        raw.insert(raw.end(), rgba + (size_t)y * width * 4, rgba + (size_t)(y + 1) * width * 4);
    }
    std::vector<uint8_t> z;  // zlib stream of stored blocks
    z.push_back(0x78), z.push_back(0x01);
    uint32_t a = 1, b = 0;  // Adler-32
    for (size_t off = 0; off < raw.size();) {
        const size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n == raw.size() ? 1 : 0);
        z.push_back(n & 0xFF), z.push_back(n >> 8), z.push_back(~n & 0xFF), z.push_back((~n >> 8) & 0xFF);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n; ++i) {
            a = (a + raw[off + i]) % 65521u;
            b = (b + a) % 65521u;
        }
        off += n;
    }
    put32(z, b << 16 | a);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    put32(ihdr, width), put32(ihdr, height);
    ihdr.push_back(8), ihdr.push_back(6), ihdr.push_back(0), ihdr.push_back(0), ihdr.push_back(0);  // 8-bit RGBA, no interlace
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});
    FILE* f = fopen(path, "wb");
    if (!f) return fail(MP_EIO, "cannot open %s for writing", path);
    const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
    fclose(f);
    return ok ? MP_OK : fail(MP_EIO, "short write to %s", path);
}

int mp_image_write_ppm(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height) {
    if (!path || !rgba || width == 0 || height == 0) return fail(MP_EINVAL, "bad argument");
    FILE* f = fopen(path, "wb");
    if (!f) return fail(MP_EIO, "cannot open %s for writing", path);
    fprintf(f, "P6\n%u %u\n255\n", width, height);
    std::vector<uint8_t> row((size_t)width * 3);
    bool ok = true;
    for (uint32_t y = 0; y < height && ok; ++y) {
        for (uint32_t x = 0; x < width; ++x) memcpy(&row[x * 3], rgba + ((size_t)y * width + x) * 4, 3);
        ok = fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    fclose(f);
    return ok ? MP_OK : fail(MP_EIO, "short write to %s", path);
}

}  // extern "C"
