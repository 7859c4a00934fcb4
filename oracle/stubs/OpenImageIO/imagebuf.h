/*
 * Stand-in for <OpenImageIO/imagebuf.h>, which is absent from this image.
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build). Declares exactly the names the
 * reference uses for image FILE I/O (meshEngine.cpp:76-88, camera.cpp:178-188);
 * none of them is on the ray-tracing path.
 *
 * ImageInput reads ONE format, PFM ("PF" = 3 channels, "Pf" = 1 channel, little or
 * big endian, rows stored bottom to top and handed out top to bottom as OpenImageIO
 * does), so that MeshEngine::bindTexture — and with it the texture arguments of the
 * reference's command line (main.cpp:71-92) — runs through the reference's own code
 * on a float texture. Every other file fails to open, as a missing file does with
 * the real library. ImageBuf::write() is a no-op (Camera::saveFrame's PNG / EXR are
 * written by the host library's own writers instead).
 */
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
namespace OpenImageIO {
struct TypeDesc {
    enum BASETYPE { UNKNOWN = 0, FLOAT = 1 };
    TypeDesc(BASETYPE = UNKNOWN) {}
};
struct ImageSpec {
    int width = 0, height = 0, nchannels = 0;
    ImageSpec(int w, int h, int c, TypeDesc = TypeDesc()) : width(w), height(h), nchannels(c) {}
};
struct ImageBuf {
    ImageBuf(const ImageSpec&, void*) {}
    bool write(const std::string&, const std::string&) { return false; }
};
struct ImageInput {
    ImageSpec m_spec{0, 0, 0};
    std::vector<float> m_pixels; // top row first

    static ImageInput* open(const std::string& name) {
        FILE* f = name.empty() ? nullptr : fopen(name.c_str(), "rb");
        if (!f) return nullptr;
        char magic[3] = {0, 0, 0};
        int w = 0, h = 0;
        float scale = 0.f;
        ImageInput* in = nullptr;
        if (fscanf(f, "%2s %d %d %f", magic, &w, &h, &scale) == 4 && fgetc(f) != EOF && w > 0 && h > 0 && scale != 0.f &&
            (!strcmp(magic, "PF") || !strcmp(magic, "Pf"))) {
            const int c = magic[1] == 'F' ? 3 : 1;
            std::vector<float> raw((size_t)w * h * c);
            if (fread(raw.data(), sizeof(float), raw.size(), f) == raw.size()) {
                if (scale > 0.f) // big endian
                    for (float& v : raw) { uint32_t u; memcpy(&u, &v, 4); u = __builtin_bswap32(u); memcpy(&v, &u, 4); }
                in = new ImageInput();
                in->m_spec = ImageSpec(w, h, c);
                in->m_pixels.resize(raw.size());
                for (int y = 0; y < h; ++y)
                    memcpy(&in->m_pixels[(size_t)y * w * c], &raw[(size_t)(h - 1 - y) * w * c], sizeof(float) * (size_t)w * c);
            }
        }
        fclose(f);
        return in;
    }
    const ImageSpec& spec() const { return m_spec; }
    bool read_image(TypeDesc, void* dst) {
        memcpy(dst, m_pixels.data(), sizeof(float) * m_pixels.size());
        return true;
    }
    bool close() { return true; }
    static void destroy(ImageInput* p) { delete p; }
};
} // namespace OpenImageIO
namespace OIIO = OpenImageIO;
