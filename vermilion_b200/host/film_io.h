/* film_io.h — host-side film output and progressive checkpoints (C linkage, in libvermilion_host.so). No CUDA, no reference headers.
 *
 * The step after the hot path: Camera::saveFrame (reference core/camera/camera.cpp:126-194) converts the film
 * to 8-bit RGBA + float depth and hands both to OpenImageIO (name.png, name_depth.exr). OpenImageIO is not in this image, so the two
 * writers here produce the same PNG (zlib) and a PFM for the depth plane. The 8-bit conversion itself runs on the device
 * (vmn_read_image_rgba8). The reference has no checkpointing (its image lives only in Camera::mImage until main.cpp:94-95); the
 * accumulation film (sum r, sum g, sum b, n) plus the sample counter is all a FIXED-mode render needs to resume, so a checkpoint
 * is that buffer with a small header. */
#ifndef VERMILION_B200_FILM_IO_H
#define VERMILION_B200_FILM_IO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* 8-bit RGBA PNG (colour type 6, no interlace), rows top to bottom as in Camera::mImage. 0 = ok. */
int vmnh_write_png_rgba8(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height);
/* PFM ("Pf" 1 channel / "PF" 3 channels, little endian, rows bottom to top as the format demands). 0 = ok. */
int vmnh_write_pfm(const char* path, const float* data, uint32_t width, uint32_t height, uint32_t channels);
/* Camera::saveFrame for an RGBAZ film already converted by vmn_read_image_rgba8: <name>.png and <name>_depth.pfm. 0 = ok. */
int vmnh_save_frame(const char* name, const uint8_t* rgba, const float* depth, uint32_t width, uint32_t height);

typedef struct {
    char magic[8];        /* "VMNFILM2" */
    uint32_t width, height;
    uint32_t spp_done;    /* samples [0, spp_done) of the sequence are in the film */
    uint32_t spp_total;   /* the camera's uSamplesPerPixel (fixes the stratum length spp/4) */
    uint64_t seed;        /* Philox key */
    uint32_t max_segments;
    uint32_t crc32;       /* of the payload */
    uint64_t frame_fingerprint; /* what the film is a picture OF: camera pose and film geometry, estimator, scene and texture content
                                 * (vmnh_hash64 chain computed by the caller, b200_pathtracer.cpp); a resume requires equality */
} vmnh_checkpoint_header;

/* 64-bit content hash (multiply-xorshift over 8-byte words, seeded; not cryptographic): scene cache key and checkpoint fingerprint */
uint64_t vmnh_hash64(const void* data, uint64_t bytes, uint64_t seed);

/* film_sum: width*height*4 floats (VMN_FILM_SUM layout). Written to <path>.tmp and renamed, so a crash never leaves a torn file. */
int vmnh_checkpoint_save(const char* path, const float* film_sum, const vmnh_checkpoint_header* hdr);
/* reads and validates the header (magic, size); 0 = ok, 1 = cannot open, 2 = not a checkpoint / truncated */
int vmnh_checkpoint_info(const char* path, vmnh_checkpoint_header* out);
/* loads the payload (n_floats must be width*height*4); 3 = size mismatch, 4 = CRC mismatch */
int vmnh_checkpoint_load(const char* path, float* film_sum, uint64_t n_floats, vmnh_checkpoint_header* out);

#ifdef __cplusplus
}
#endif
#endif
