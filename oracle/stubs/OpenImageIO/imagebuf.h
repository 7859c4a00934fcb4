/*
 * Stand-in for <OpenImageIO/imagebuf.h>, which is absent from this image.
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build). Declares exactly the names the
 * reference uses for image FILE I/O (meshEngine.cpp:76-88, camera.cpp:178-188);
 * none of them is on the ray-tracing path. open() fails, write() is a no-op.
 */
#pragma once
#include <string>
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
    static ImageInput* open(const std::string&) { return nullptr; }
    const ImageSpec& spec() const { static ImageSpec s(0, 0, 0); return s; }
    bool read_image(TypeDesc, void*) { return false; }
    bool close() { return true; }
    static void destroy(ImageInput*) {}
};
} // namespace OpenImageIO
namespace OIIO = OpenImageIO;
