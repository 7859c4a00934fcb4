/*
 * scenes.cpp — procedural scene generators for the configurations BASELINE.json
 * names (SURVEY.md §8d C1–C4). Host C++; no CUDA. The same arrays feed the GPU
 * engine (vmn_scene_set_triangles) and the CPU checker, so both sides see
 * bit-identical geometry.
 *
 * Triangle arrays use the layout MeshEngine::createBVH produces per face
 * (reference core/engines/meshEngine.cpp:660-717): 9 floats of positions
 * (v0,v1,v2), 9 floats of vertex normals, 6 floats of uv per triangle.
 * All generators are deterministic; the seeded ones use std::mt19937 +
 * uniform_real_distribution<float> (libstdc++), as SURVEY.md §8d C3 specifies.
 */
#include "scenes.h"

#include <cmath>
#include <cstring>
#include <map>
#include <random>
#include <vector>
#include <cstdio>

namespace {

struct P3 { float x, y, z; };
inline P3 operator+(P3 a, P3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline P3 operator-(P3 a, P3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline P3 operator*(P3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline P3 cross(P3 a, P3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline P3 norm(P3 a) { float l = std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); return l > 0 ? P3{a.x / l, a.y / l, a.z / l} : a; }

struct Sink {
    float *pos, *nrm, *uv;
    uint64_t n = 0;
    void tri(P3 a, P3 b, P3 c, P3 na, P3 nb, P3 nc) {
        if (pos) { float* p = pos + 9 * n; p[0] = a.x; p[1] = a.y; p[2] = a.z; p[3] = b.x; p[4] = b.y; p[5] = b.z; p[6] = c.x; p[7] = c.y; p[8] = c.z; }
        if (nrm) { float* q = nrm + 9 * n; q[0] = na.x; q[1] = na.y; q[2] = na.z; q[3] = nb.x; q[4] = nb.y; q[5] = nb.z; q[6] = nc.x; q[7] = nc.y; q[8] = nc.z; }
        if (uv) { float* t = uv + 6 * n; t[0] = 0; t[1] = 0; t[2] = 1; t[3] = 0; t[4] = 0; t[5] = 1; }
        ++n;
    }
    void flat(P3 a, P3 b, P3 c) { P3 fn = norm(cross(b - a, c - a)); tri(a, b, c, fn, fn, fn); }
    void quad(P3 a, P3 b, P3 c, P3 d) { flat(a, b, c); flat(a, c, d); }
};

// unit icosphere, `level` 4-to-1 subdivisions: 20 * 4^level triangles
void icosphere(int level, std::vector<P3>& verts, std::vector<uint32_t>& idx) {
    const float t = (1.f + std::sqrt(5.f)) * 0.5f;
    P3 base[12] = {{-1, t, 0}, {1, t, 0}, {-1, -t, 0}, {1, -t, 0}, {0, -1, t}, {0, 1, t},
                   {0, -1, -t}, {0, 1, -t}, {t, 0, -1}, {t, 0, 1}, {-t, 0, -1}, {-t, 0, 1}};
    static const uint32_t faces[60] = {0, 11, 5, 0, 5, 1, 0, 1, 7, 0, 7, 10, 0, 10, 11, 1, 5, 9, 5, 11, 4, 11, 10, 2, 10, 7, 6, 7, 1, 8,
                                       3, 9, 4, 3, 4, 2, 3, 2, 6, 3, 6, 8, 3, 8, 9, 4, 9, 5, 2, 4, 11, 6, 2, 10, 8, 6, 7, 9, 8, 1};
    verts.clear(); idx.assign(faces, faces + 60);
    for (auto& b : base) verts.push_back(norm(b));
    for (int l = 0; l < level; ++l) {
        std::map<uint64_t, uint32_t> mid;
        auto midpoint = [&](uint32_t a, uint32_t b) {
            uint64_t key = a < b ? ((uint64_t)a << 32) | b : ((uint64_t)b << 32) | a;
            auto it = mid.find(key);
            if (it != mid.end()) return it->second;
            verts.push_back(norm((verts[a] + verts[b]) * 0.5f));
            return mid[key] = (uint32_t)verts.size() - 1;
        };
        std::vector<uint32_t> next; next.reserve(idx.size() * 4);
        for (size_t f = 0; f < idx.size(); f += 3) {
            uint32_t a = idx[f], b = idx[f + 1], c = idx[f + 2];
            uint32_t ab = midpoint(a, b), bc = midpoint(b, c), ca = midpoint(c, a);
            uint32_t tri[12] = {a, ab, ca, b, bc, ab, c, ca, bc, ab, bc, ca};
            next.insert(next.end(), tri, tri + 12);
        }
        idx.swap(next);
    }
}

} // namespace

extern "C" {

/* C1: Cornell box, 30 triangles: floor, ceiling, back, left, right (2 each) + a short and a
 * tall block with 5 faces each (no bottom). Sits on the reference's hard-coded floor y = 0,
 * open towards +z, around emitter A (15,140,25) (meshEngine.cpp:377). */
uint64_t vmn_gen_cornell(float* pos, float* nrm, float* uv) {
    Sink s{pos, nrm, uv};
    const float X = 275.f, Y = 550.f, Z = 275.f;
    s.quad({-X, 0, -Z}, {-X, 0, Z}, {X, 0, Z}, {X, 0, -Z});     // floor
    s.quad({-X, Y, -Z}, {X, Y, -Z}, {X, Y, Z}, {-X, Y, Z});     // ceiling
    s.quad({-X, 0, -Z}, {X, 0, -Z}, {X, Y, -Z}, {-X, Y, -Z});   // back
    s.quad({-X, 0, -Z}, {-X, Y, -Z}, {-X, Y, Z}, {-X, 0, Z});   // left
    s.quad({X, 0, -Z}, {X, 0, Z}, {X, Y, Z}, {X, Y, -Z});       // right
    auto block = [&](float cx, float cz, float hw, float h, float ang) {
        float c = std::cos(ang), sn = std::sin(ang);
        P3 b[4], t[4];
        const float ox[4] = {-hw, hw, hw, -hw}, oz[4] = {-hw, -hw, hw, hw};
        for (int i = 0; i < 4; ++i) {
            float x = cx + ox[i] * c - oz[i] * sn, z = cz + ox[i] * sn + oz[i] * c;
            b[i] = {x, 0.5f, z}; t[i] = {x, h, z};
        }
        s.quad(t[0], t[3], t[2], t[1]); // top
        for (int i = 0; i < 4; ++i) { int j = (i + 1) & 3; s.quad(b[i], t[i], t[j], b[j]); }
    };
    block(90.f, 80.f, 80.f, 165.f, 0.29f);    // short block
    block(-95.f, -90.f, 80.f, 330.f, -0.31f); // tall block
    return s.n;
}

/* C2: nx*ny icospheres of subdivision `level` (20*4^level triangles each), radius r, centres on
 * a fixed grid inside the reference room (walls x,z = ±2000, floor 0, ceiling 1000). Seed-free. */
uint64_t vmn_gen_spheres_count(uint32_t nx, uint32_t ny, uint32_t level) {
    uint64_t per = 20; for (uint32_t l = 0; l < level; ++l) per *= 4; return per * nx * ny;
}
uint64_t vmn_gen_spheres(float* pos, float* nrm, float* uv, uint32_t nx, uint32_t ny, uint32_t level, float radius) {
    std::vector<P3> verts; std::vector<uint32_t> idx;
    icosphere((int)level, verts, idx);
    Sink s{pos, nrm, uv};
    for (uint32_t j = 0; j < ny; ++j)
        for (uint32_t i = 0; i < nx; ++i) {
            P3 c{(nx > 1 ? (-1.f + 2.f * i / (nx - 1)) : 0.f) * 1200.f,
                 180.f + (ny > 1 ? (float)j / (ny - 1) : 0.5f) * 640.f,
                 -600.f + 350.f * (float)((i + j) & 1)};
            for (size_t f = 0; f < idx.size(); f += 3) {
                P3 a = verts[idx[f]], b = verts[idx[f + 1]], d = verts[idx[f + 2]];
                s.tri(c + a * radius, c + b * radius, c + d * radius, a, b, d);
            }
        }
    return s.n;
}

/* C3 soup: n triangles, centres uniform in [0,1000]^3, vertices = centre + U[-1/2,1/2]^3 * 2*1000/cbrt(n);
 * std::mt19937(seed), uniform_real_distribution<float>. Normals = face normal. */
uint64_t vmn_gen_soup(float* pos, float* nrm, float* uv, uint64_t n, uint32_t seed) {
    std::mt19937 eng(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    const float ext = 2.f * 1000.f / std::cbrt((float)n);
    Sink s{pos, nrm, uv};
    for (uint64_t i = 0; i < n; ++i) {
        P3 c{U(eng) * 1000.f, U(eng) * 1000.f, U(eng) * 1000.f};
        P3 v[3];
        for (int k = 0; k < 3; ++k) v[k] = {c.x + (U(eng) - 0.5f) * ext, c.y + (U(eng) - 0.5f) * ext, c.z + (U(eng) - 0.5f) * ext};
        s.flat(v[0], v[1], v[2]);
    }
    return s.n;
}

/* C3 rays: origin uniform in [lo,hi]^3, direction uniform on the sphere; 8 floats per ray
 * (o.xyz, 0, d.xyz, 0) = vmn_ray. */
void vmn_gen_rays_uniform(float* rays, uint64_t n, uint32_t seed, float lo, float hi) {
    std::mt19937 eng(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    for (uint64_t i = 0; i < n; ++i) {
        float* r = rays + 8 * i;
        r[0] = lo + U(eng) * (hi - lo); r[1] = lo + U(eng) * (hi - lo); r[2] = lo + U(eng) * (hi - lo); r[3] = 0.f;
        float z = 2.f * U(eng) - 1.f, phi = 6.2831853071795864769f * U(eng);
        float rr = std::sqrt(std::fmax(0.f, 1.f - z * z));
        r[4] = rr * std::cos(phi); r[5] = rr * std::sin(phi); r[6] = z; r[7] = 0.f;
    }
}

/* C4 city: gx*gz blocks on the floor, each one of 8 procedurally tessellated tower meshes
 * (a box column subdivided `tess` times per side, with a seeded facade displacement),
 * per-instance transform baked in (the reference flattens instances, meshEngine.cpp:106). */
static uint64_t city_building(Sink& s, float cx, float cz, float hw, float h, uint32_t tess, uint32_t kind, std::mt19937& eng) {
    uint64_t before = s.n;
    std::uniform_real_distribution<float> U(0.f, 1.f);
    const float bump = hw * 0.06f * (1.f + (kind & 3));
    const float phase = U(eng) * 6.28f;
    const int nu = (int)tess, nv = (int)tess * (2 + (kind >> 2));
    // four facades
    for (int side = 0; side < 4; ++side) {
        float ax = (side == 0) ? 1.f : (side == 1) ? 0.f : (side == 2) ? -1.f : 0.f; // tangent
        float az = (side == 0) ? 0.f : (side == 1) ? 1.f : (side == 2) ? 0.f : -1.f;
        float nx = az, nz = -ax; // outward normal
        auto P = [&](int iu, int iv) {
            float u = (float)iu / nu * 2.f - 1.f, v = (float)iv / nv;
            float d = (iu == 0 || iu == nu) ? 0.f : bump * std::sin(phase + 9.f * u + 14.f * v * (1 + (kind & 1)));
            return P3{cx + ax * u * hw + nx * (hw + d), v * h, cz + az * u * hw + nz * (hw + d)};
        };
        for (int iv = 0; iv < nv; ++iv)
            for (int iu = 0; iu < nu; ++iu) s.quad(P(iu, iv), P(iu, iv + 1), P(iu + 1, iv + 1), P(iu + 1, iv));
    }
    // roof
    for (int iz = 0; iz < nu; ++iz)
        for (int ix = 0; ix < nu; ++ix) {
            auto R = [&](int a, int b) { return P3{cx + ((float)a / nu * 2.f - 1.f) * hw, h, cz + ((float)b / nu * 2.f - 1.f) * hw}; };
            s.quad(R(ix, iz), R(ix, iz + 1), R(ix + 1, iz + 1), R(ix + 1, iz));
        }
    return s.n - before;
}
uint64_t vmn_gen_city_count(uint32_t gx, uint32_t gz, uint32_t tess, uint32_t seed) {
    Sink s{nullptr, nullptr, nullptr};
    std::mt19937 eng(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    for (uint32_t j = 0; j < gz; ++j)
        for (uint32_t i = 0; i < gx; ++i) { uint32_t kind = (uint32_t)(U(eng) * 8.f) & 7u; float h = U(eng); (void)h; city_building(s, 0, 0, 1, 1, tess, kind, eng); }
    return s.n;
}
uint64_t vmn_gen_city(float* pos, float* nrm, float* uv, uint32_t gx, uint32_t gz, uint32_t tess, uint32_t seed) {
    Sink s{pos, nrm, uv};
    std::mt19937 eng(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    const float span = 3600.f, pitchx = span / gx, pitchz = span / gz;
    for (uint32_t j = 0; j < gz; ++j)
        for (uint32_t i = 0; i < gx; ++i) {
            uint32_t kind = (uint32_t)(U(eng) * 8.f) & 7u;
            float h = 60.f + U(eng) * 700.f;
            float cx = -span * 0.5f + (i + 0.5f) * pitchx, cz = -span * 0.5f + (j + 0.5f) * pitchz;
            city_building(s, cx, cz, 0.36f * std::fmin(pitchx, pitchz), h, tess, kind, eng);
        }
    return s.n;
}

} // extern "C"
