/* scenes.h — procedural scene generators (host C++, C linkage). See scenes.cpp. */
#ifndef VERMILION_B200_SCENES_H
#define VERMILION_B200_SCENES_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* Every generator writes 9 floats (positions), 9 floats (normals), 6 floats (uv) per triangle;
 * any of the three output pointers may be NULL. Returns the number of triangles. */
uint64_t vmn_gen_cornell(float* pos, float* nrm, float* uv); /* 30 triangles */
uint64_t vmn_gen_spheres_count(uint32_t nx, uint32_t ny, uint32_t level);
uint64_t vmn_gen_spheres(float* pos, float* nrm, float* uv, uint32_t nx, uint32_t ny, uint32_t level, float radius);
uint64_t vmn_gen_soup(float* pos, float* nrm, float* uv, uint64_t n, uint32_t seed);
void vmn_gen_rays_uniform(float* rays8, uint64_t n, uint32_t seed, float lo, float hi);
uint64_t vmn_gen_city_count(uint32_t gx, uint32_t gz, uint32_t tess, uint32_t seed);
uint64_t vmn_gen_city(float* pos, float* nrm, float* uv, uint32_t gx, uint32_t gz, uint32_t tess, uint32_t seed);
#ifdef __cplusplus
}
#endif
#endif
