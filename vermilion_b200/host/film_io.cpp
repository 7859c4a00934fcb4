// film_io.cpp — see film_io.h. PNG through zlib (deflate + crc32), PFM, and the progressive-render checkpoint file.
#include "film_io.h"

#include <zlib.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

void put_be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back(uint8_t(x >> 24)); v.push_back(uint8_t(x >> 16)); v.push_back(uint8_t(x >> 8)); v.push_back(uint8_t(x));
}

void png_chunk(std::vector<uint8_t>& out, const char type[4], const uint8_t* data, size_t n) {
    put_be32(out, (uint32_t)n);
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    put_be32(out, (uint32_t)crc32(0L, out.data() + start, (uInt)(n + 4)));
}

bool write_file(const char* path, const void* data, size_t n) {
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    const bool ok = fwrite(data, 1, n, f) == n;
    return (fclose(f) == 0) && ok;
}

} // namespace

extern "C" {

int vmnh_write_png_rgba8(const char* path, const uint8_t* rgba, uint32_t w, uint32_t h) {
    if (!path || !rgba || !w || !h) return 1;
    // filter type 0 in front of every scanline
    std::vector<uint8_t> raw((size_t)h * (4 * (size_t)w + 1));
    for (uint32_t y = 0; y < h; ++y) {
        uint8_t* row = raw.data() + (size_t)y * (4 * (size_t)w + 1);
        row[0] = 0;
        memcpy(row + 1, rgba + (size_t)y * 4 * w, 4 * (size_t)w);
    }
    uLongf zn = compressBound((uLong)raw.size());
    std::vector<uint8_t> z(zn);
    if (compress2(z.data(), &zn, raw.data(), (uLong)raw.size(), 6) != Z_OK) return 2;
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, w); put_be32(ihdr, h);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0); // 8 bit, RGBA, deflate, no filter set, no interlace
    png_chunk(out, "IHDR", ihdr.data(), ihdr.size());
    png_chunk(out, "IDAT", z.data(), zn);
    png_chunk(out, "IEND", nullptr, 0);
    return write_file(path, out.data(), out.size()) ? 0 : 3;
}

int vmnh_write_pfm(const char* path, const float* data, uint32_t w, uint32_t h, uint32_t channels) {
    if (!path || !data || !w || !h || (channels != 1 && channels != 3)) return 1;
    FILE* f = fopen(path, "wb");
    if (!f) return 3;
    fprintf(f, "%s\n%u %u\n-1.0\n", channels == 1 ? "Pf" : "PF", w, h); // negative scale = little endian
    bool ok = true;
    for (uint32_t y = h; y-- > 0 && ok;) ok = fwrite(data + (size_t)y * w * channels, sizeof(float), (size_t)w * channels, f) == (size_t)w * channels;
    return (fclose(f) == 0 && ok) ? 0 : 3;
}

int vmnh_save_frame(const char* name, const uint8_t* rgba, const float* depth, uint32_t w, uint32_t h) {
    if (!name) return 1;
    const std::string base(name);
    int rc = vmnh_write_png_rgba8((base + ".png").c_str(), rgba, w, h); // camera.cpp:178-181
    if (rc == 0 && depth) rc = vmnh_write_pfm((base + "_depth.pfm").c_str(), depth, w, h, 1); // camera.cpp:184-189 writes an EXR there
    return rc;
}

uint64_t vmnh_hash64(const void* data, uint64_t bytes, uint64_t seed) {
    const unsigned char* p = static_cast<const unsigned char*>(data);
    uint64_t h = seed ^ (bytes * 0x9E3779B97F4A7C15ull);
    auto mix = [&h](uint64_t w) { h ^= w; h *= 0xFF51AFD7ED558CCDull; h ^= h >> 32; };
    uint64_t i = 0;
    for (; i + 8 <= bytes; i += 8) { uint64_t w; memcpy(&w, p + i, 8); mix(w); }
    if (i < bytes) { uint64_t w = 0; memcpy(&w, p + i, bytes - i); mix(w); }
    h ^= h >> 29; h *= 0xC4CEB9FE1A85EC53ull; h ^= h >> 32;
    return h;
}

int vmnh_checkpoint_save(const char* path, const float* film, const vmnh_checkpoint_header* hdr) {
    if (!path || !film || !hdr || !hdr->width || !hdr->height) return 1;
    vmnh_checkpoint_header h = *hdr;
    memcpy(h.magic, "VMNFILM2", 8);
    const size_t bytes = (size_t)h.width * h.height * 16;
    uLong crc = crc32(0L, Z_NULL, 0);
    for (size_t off = 0; off < bytes; off += (1u << 30)) { // zlib takes 32-bit lengths
        const size_t n = bytes - off < (1u << 30) ? bytes - off : (1u << 30);
        crc = crc32(crc, reinterpret_cast<const Bytef*>(film) + off, (uInt)n);
    }
    h.crc32 = (uint32_t)crc;
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return 3;
    bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(film, 1, bytes, f) == bytes;
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return 3; }
    return 0;
}

int vmnh_checkpoint_info(const char* path, vmnh_checkpoint_header* out) {
    if (!path || !out) return 1;
    FILE* f = fopen(path, "rb");
    if (!f) return 1;
    vmnh_checkpoint_header h;
    const bool ok = fread(&h, sizeof(h), 1, f) == 1;
    long size = -1;
    if (ok && fseek(f, 0, SEEK_END) == 0) size = ftell(f);
    fclose(f);
    if (!ok || memcmp(h.magic, "VMNFILM2", 8) != 0) return 2;
    if (size < 0 || (uint64_t)size != sizeof(h) + (uint64_t)h.width * h.height * 16) return 2;
    *out = h;
    return 0;
}

int vmnh_checkpoint_load(const char* path, float* film, uint64_t n_floats, vmnh_checkpoint_header* out) {
    vmnh_checkpoint_header h;
    int rc = vmnh_checkpoint_info(path, &h);
    if (rc != 0) return rc;
    if (!film || n_floats != (uint64_t)h.width * h.height * 4) return 3;
    FILE* f = fopen(path, "rb");
    if (!f) return 1;
    const size_t bytes = (size_t)n_floats * 4;
    const bool ok = fseek(f, (long)sizeof(h), SEEK_SET) == 0 && fread(film, 1, bytes, f) == bytes;
    fclose(f);
    if (!ok) return 2;
    uLong crc = crc32(0L, Z_NULL, 0);
    for (size_t off = 0; off < bytes; off += (1u << 30)) {
        const size_t n = bytes - off < (1u << 30) ? bytes - off : (1u << 30);
        crc = crc32(crc, reinterpret_cast<const Bytef*>(film) + off, (uInt)n);
    }
    if ((uint32_t)crc != h.crc32) return 4;
    if (out) *out = h;
    return 0;
}

} // extern "C"
