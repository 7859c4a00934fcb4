"""ctypes bindings for the CPU checker (oracle/) — TEST INFRASTRUCTURE ONLY.

`Oracle`  : oracle/liboracle.so      — our restatement (always available; built by `make -C oracle oracle`).
`Ref`     : oracle/_ref/libvermref.so — the reference's own sources compiled unmodified (built here from
            /root/reference by `make -C oracle ref`; the prebuilt .so travels to the GPU box).
Nothing under vermilion_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

RAY = np.dtype([("o", "<f4", 3), ("pad0", "<f4"), ("d", "<f4", 3), ("pad1", "<f4")])
HIT = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<u4")])
RAYCAST = np.dtype([("P", "<f4", 3), ("dist", "<f4"), ("N", "<f4", 3), ("flags", "<u4"), ("uv", "<f4", 2),
                    ("Le", "<f4", 3), ("tri", "<u4"), ("pad", "<u4", 2)])
assert RAY.itemsize == 32 and HIT.itemsize == 16 and RAYCAST.itemsize == 64
RC_HIT, RC_MATERIAL = 1, 2
MODE_COMPAT, MODE_FIXED = 0, 1
RNG_MT64, RNG_PHILOX, RNG_MT32_SHARED = 0, 1, 2
BF_SHADOW = 1
MISS = 0xFFFFFFFF


class OrcCamera(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("position", C.c_float * 3), ("rotation", C.c_float * 3),
                ("dist_to_film", C.c_float), ("sensor_x", C.c_float), ("sensor_y", C.c_float), ("spp", C.c_uint32)]


class RefCameraSettings(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("position", C.c_float * 3), ("rotation_deg", C.c_float * 3),
                ("back_distance", C.c_float), ("back_size_x", C.c_float), ("back_size_y", C.c_float), ("rays_per_pixel", C.c_uint32)]


def camera_settings(width, height, position, rotation_deg, back_distance=6.0, back_size=(3.6, 2.4), spp=16):
    return RefCameraSettings(width, height, (C.c_float * 3)(*position), (C.c_float * 3)(*rotation_deg), back_distance,
                             back_size[0], back_size[1], spp)


def camera_fields_py(cs):
    """Vermilion::Camera ctor (camera.cpp:34-56) restated: degrees → radians in double, x,y negated, narrowed to float."""
    cam = OrcCamera()
    cam.width, cam.height = cs.width, cs.height
    for i in range(3):
        cam.position[i] = cs.position[i]
    rot = [-float(cs.rotation_deg[0]) * 3.1415926535 / 180, -float(cs.rotation_deg[1]) * 3.1415926535 / 180,
           float(cs.rotation_deg[2]) * 3.1415926535 / 180]
    for i in range(3):
        cam.rotation[i] = np.float32(rot[i])
    cam.dist_to_film = cs.back_distance
    cam.sensor_x, cam.sensor_y = cs.back_size_x, cs.back_size_y
    cam.spp = cs.rays_per_pixel
    return cam


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t) if a is not None else None


def make_rays(o, d):
    o = np.asarray(o, np.float32).reshape(-1, 3)
    d = np.asarray(d, np.float32).reshape(-1, 3)
    r = np.zeros(len(o), RAY)
    r["o"], r["d"] = o, d
    return r


def build_oracle(quiet=True):
    """(re)build oracle/liboracle.so and, when /root/reference exists, oracle/_ref/*.so"""
    subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None, stderr=subprocess.STDOUT if quiet else None)


class Oracle:
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            path = os.path.join(ORACLE_DIR, "liboracle.so")
            if not os.path.exists(path):
                build_oracle()
            L = C.CDLL(path)
            L.orc_scene_create.restype = C.c_void_p
            L.orc_scene_create.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]
            L.orc_scene_destroy.argtypes = [C.c_void_p]
            L.orc_scene_set_texture.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
            L.orc_scene_set_spheres.argtypes = [C.c_void_p, C.c_int]
            L.orc_scene_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
            L.orc_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
            L.orc_intersect_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
            L.orc_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
            L.orc_raycast_many.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
            L.orc_radiance_many.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
            L.orc_camera_matrix.argtypes = [C.POINTER(OrcCamera), C.c_void_p]
            L.orc_camera_ray.argtypes = [C.POINTER(OrcCamera), C.c_uint64, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_void_p]
            L.orc_render.restype = C.c_uint64
            L.orc_render.argtypes = [C.c_void_p, C.POINTER(OrcCamera), C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64,
                                     C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
            L.orc_scene_add_texture.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
            L.orc_render_bruteforce.restype = C.c_uint64
            L.orc_render_bruteforce.argtypes = [C.c_void_p, C.POINTER(OrcCamera), C.c_int, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64,
                                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
            L.orc_philox.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]
            L.orc_max_threads.restype = C.c_int
            cls._lib = L
        return cls._lib

    def __init__(self, pos, nrm=None, uv=None, leaf_size=4, spheres=True):
        L = self.lib()
        self.pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 9)
        self.nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float32).reshape(-1, 9)
        self.uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 6)
        self.h = L.orc_scene_create(_p(self.pos), _p(self.nrm), _p(self.uv), len(self.pos), leaf_size)
        L.orc_scene_set_spheres(self.h, int(spheres))

    def close(self):
        if self.h:
            self.lib().orc_scene_destroy(self.h)
            self.h = None

    __del__ = close

    def set_texture(self, tex):
        tex = np.ascontiguousarray(tex, np.float32)
        h, w, c = tex.shape
        self.lib().orc_scene_set_texture(self.h, _p(tex), w, h, c)

    def add_texture(self, tex):
        tex = np.ascontiguousarray(tex, np.float32)
        h, w, c = tex.shape
        self.lib().orc_scene_add_texture(self.h, _p(tex), w, h, c)

    def render_bruteforce(self, cam, rng=RNG_PHILOX, seed=1, flags=0, p0=0, p1=None, threads=0, want_raw=False):
        """BruteForceTracer::Render restated: (image[npix,5], samples_taken[npix], BVH queries[, raw[npix,6]])"""
        npx = cam.width * cam.height
        p1 = npx if p1 is None else p1
        img = np.zeros((npx, 5), np.float32)
        taken = np.zeros(npx, np.uint32)
        raw = np.zeros((npx, 6), np.float32) if want_raw else None
        n = self.lib().orc_render_bruteforce(self.h, C.byref(cam), rng, seed, flags, p0, p1, _p(img), _p(taken), _p(raw), threads)
        return (img, taken, n, raw) if want_raw else (img, taken, n)

    def stats(self):
        a, b = C.c_uint32(), C.c_uint32()
        self.lib().orc_scene_stats(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def intersect(self, rays, threads=0):
        out = np.zeros(len(rays), HIT)
        self.lib().orc_intersect(self.h, _p(rays), len(rays), _p(out), threads)
        return out

    def intersect_stats(self, rays):
        a, b = C.c_uint64(), C.c_uint64()
        self.lib().orc_intersect_stats(self.h, _p(rays), len(rays), C.byref(a), C.byref(b))
        return a.value, b.value

    def occluded(self, rays, threads=0):
        out = np.zeros(len(rays), np.uint8)
        self.lib().orc_occluded(self.h, _p(rays), len(rays), _p(out), threads)
        return out

    def raycast(self, rays, threads=0):
        out = np.zeros(len(rays), RAYCAST)
        self.lib().orc_raycast_many(self.h, _p(rays), len(rays), _p(out), threads)
        return out

    def radiance(self, rays, seed, rng=RNG_MT64, mode=MODE_COMPAT, threads=0):
        out = np.zeros((len(rays), 4), np.float32)
        cnt = np.zeros(len(rays), np.uint32)
        self.lib().orc_radiance_many(self.h, _p(rays), len(rays), seed, rng, mode, _p(out), _p(cnt), threads)
        return out, cnt

    def render(self, cam, mode=MODE_COMPAT, rng=RNG_PHILOX, seed=1, p0=0, p1=None, s0=0, s1=None, max_segments=5, threads=0,
               want_raw=False):
        npx = cam.width * cam.height
        p1 = npx if p1 is None else p1
        s1 = cam.spp if s1 is None else s1
        img = np.zeros((npx, 5), np.float32)
        raw = np.zeros((npx, 4), np.float32) if want_raw else None
        nrays = self.lib().orc_render(self.h, C.byref(cam), mode, rng, seed, p0, p1, s0, s1, max_segments, _p(img), _p(raw), threads)
        return (img, raw, nrays) if want_raw else (img, nrays)

    @classmethod
    def camera_matrix(cls, cam):
        m = np.zeros(16, np.float32)
        cls.lib().orc_camera_matrix(C.byref(cam), _p(m))
        return m

    @classmethod
    def camera_ray(cls, cam, p, sx, sy, jx, jy):
        r = np.zeros(1, RAY)
        cls.lib().orc_camera_ray(C.byref(cam), p, sx, sy, jx, jy, _p(r))
        return r

    @classmethod
    def philox(cls, seed, pixel, sample, block):
        out = np.zeros(4, np.uint32)
        cls.lib().orc_philox(seed, pixel, sample, block, _p(out))
        return out

    @classmethod
    def max_threads(cls):
        return cls.lib().orc_max_threads()


class Ref:
    """The reference's own compiled code. variant 'ieee' (-O2 -ffp-contract=off: THE parity oracle) or 'ofast' (as shipped)."""
    _libs = {}

    @staticmethod
    def path(variant="ieee"):
        return os.path.join(ORACLE_DIR, "_ref", "libvermref.so" if variant == "ieee" else "libvermref_ofast.so")

    @classmethod
    def available(cls, variant="ieee"):
        return os.path.exists(cls.path(variant))

    @classmethod
    def lib(cls, variant="ieee"):
        if variant not in cls._libs:
            L = C.CDLL(cls.path(variant))
            L.ref_scene_create.restype = C.c_void_p
            L.ref_scene_create.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]
            L.ref_scene_destroy.argtypes = [C.c_void_p]
            L.ref_scene_set_texture.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
            L.ref_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
            L.ref_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
            L.ref_raycast_many.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
            L.ref_radiance_many.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p, C.c_int]
            L.ref_camera_fields.argtypes = [C.POINTER(RefCameraSettings), C.POINTER(OrcCamera)]
            L.ref_camera_ray.argtypes = [C.POINTER(RefCameraSettings), C.c_uint64, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_void_p]
            L.ref_render_pathtracer.restype = C.c_double
            L.ref_render_pathtracer.argtypes = [C.c_void_p, C.POINTER(RefCameraSettings), C.c_void_p, C.c_int]
            L.ref_render_seeded.argtypes = [C.c_void_p, C.POINTER(RefCameraSettings), C.c_int, C.c_uint64, C.c_uint64, C.c_uint64,
                                            C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
            L.ref_render_bruteforce.restype = C.c_double
            L.ref_render_bruteforce.argtypes = [C.c_void_p, C.POINTER(RefCameraSettings), C.c_uint64, C.c_void_p, C.c_int]
            L.ref_scene_bind_texture_file.argtypes = [C.c_void_p, C.c_char_p]
            L.ref_scene_load_file.restype = C.c_void_p
            L.ref_scene_load_file.argtypes = [C.c_char_p]
            L.ref_scene_triangle_count.restype = C.c_uint64
            L.ref_scene_triangle_count.argtypes = [C.c_void_p]
            L.ref_scene_mesh_count.restype = C.c_uint32
            L.ref_scene_mesh_count.argtypes = [C.c_void_p]
            L.ref_scene_dump_triangles.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
            L.ref_build_flags.restype = C.c_char_p
            L.ref_max_threads.restype = C.c_int
            cls._libs[variant] = L
        return cls._libs[variant]

    def __init__(self, pos, nrm=None, uv=None, variant="ieee", mesh_engine=True):
        self.L = self.lib(variant)
        self.pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 9)
        self.nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float32).reshape(-1, 9)
        self.uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 6)
        self.h = self.L.ref_scene_create(_p(self.pos), _p(self.nrm), _p(self.uv), len(self.pos), int(mesh_engine))

    @classmethod
    def load_file(cls, path, variant="ieee"):
        """The reference's own loading path: MeshEngine::load (its flags, meshEngine.cpp:103-125) through the vendored Assimp.
        pos / nrm / uv are what MeshEngine::createBVH's walk read (meshEngine.cpp:660-717) — the canonical triangle order."""
        self = cls.__new__(cls)
        self.L = cls.lib(variant)
        self.h = self.L.ref_scene_load_file(os.fsencode(path))
        if not self.h:
            raise RuntimeError(f"MeshEngine::load({path}) failed")
        n = self.L.ref_scene_triangle_count(self.h)
        self.n_meshes = self.L.ref_scene_mesh_count(self.h)
        self.pos, self.nrm, self.uv = np.empty((n, 9), np.float32), np.empty((n, 9), np.float32), np.empty((n, 6), np.float32)
        self.L.ref_scene_dump_triangles(self.h, _p(self.pos), _p(self.nrm), _p(self.uv))
        return self

    def close(self):
        if getattr(self, "h", None):
            self.L.ref_scene_destroy(self.h)
            self.h = None

    __del__ = close

    def set_texture(self, tex):
        tex = np.ascontiguousarray(tex, np.float32)
        h, w, c = tex.shape
        self.L.ref_scene_set_texture(self.h, _p(tex), w, h, c)

    def bind_texture_file(self, path):
        """MeshEngine::bindTexture(path) — the reference's own loader (meshEngine.cpp:74-93); the oracle build reads PFM."""
        return bool(self.L.ref_scene_bind_texture_file(self.h, os.fsencode(path)))

    def intersect(self, rays, threads=0):
        out = np.zeros(len(rays), HIT)
        self.L.ref_intersect(self.h, _p(rays), len(rays), _p(out), threads)
        return out

    def occluded(self, rays, threads=0):
        out = np.zeros(len(rays), np.uint8)
        self.L.ref_occluded(self.h, _p(rays), len(rays), _p(out), threads)
        return out

    def raycast(self, rays, threads=0):
        out = np.zeros(len(rays), RAYCAST)
        self.L.ref_raycast_many(self.h, _p(rays), len(rays), _p(out), threads)
        return out

    def radiance(self, rays, seed, mode=MODE_COMPAT, threads=0):
        out = np.zeros((len(rays), 4), np.float32)
        self.L.ref_radiance_many(self.h, _p(rays), len(rays), seed, mode, _p(out), threads)
        return out

    def camera_fields(self, cs):
        cam = OrcCamera()
        self.L.ref_camera_fields(C.byref(cs), C.byref(cam))
        return cam

    def camera_ray(self, cs, p, sx, sy, jx, jy):
        r = np.zeros(1, RAY)
        self.L.ref_camera_ray(C.byref(cs), p, sx, sy, jx, jy, _p(r))
        return r

    def render_pathtracer(self, cs, threads=0):
        img = np.zeros((cs.width * cs.height, 5), np.float32)
        secs = self.L.ref_render_pathtracer(self.h, C.byref(cs), _p(img), threads)
        return img, secs

    def render_bruteforce(self, cs, seed=1, threads=1):
        """the real BruteForceTracer::Render with its time(0) seed pinned; threads=1 is deterministic"""
        img = np.zeros((cs.width * cs.height, 5), np.float32)
        secs = self.L.ref_render_bruteforce(self.h, C.byref(cs), seed, _p(img), threads)
        return img, secs

    def render_seeded(self, cs, mode=MODE_COMPAT, seed=1, p0=0, p1=None, s0=0, s1=None, max_segments=5, threads=0, want_raw=False):
        npx = cs.width * cs.height
        p1 = npx if p1 is None else p1
        s1 = cs.rays_per_pixel if s1 is None else s1
        img = np.zeros((npx, 5), np.float32)
        raw = np.zeros((npx, 4), np.float32) if want_raw else None
        self.L.ref_render_seeded(self.h, C.byref(cs), mode, seed, p0, p1, s0, s1, max_segments, _p(img), _p(raw), threads)
        return (img, raw) if want_raw else img
