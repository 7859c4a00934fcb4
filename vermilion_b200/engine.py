"""ctypes binding of libvermilion_b200.so (include/vermilion_b200.h) — plumbing for tests and bench.py.

The product is the C-ABI library; the reference-side binding is the C++ adapter in vermilion_b200/host/
(class B200PathTracer : Vermilion::Integrator). This module only marshals numpy arrays across the same ABI.
There is NO CPU fallback: if the library is missing, or no CUDA device is present, construction raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VMN_LIB") or os.path.join(_HERE, "libvermilion_b200.so")  # VMN_LIB: experiment variants (csrc/Makefile)

RAY = np.dtype([("o", "<f4", 3), ("pad0", "<f4"), ("d", "<f4", 3), ("pad1", "<f4")])
HIT = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<u4")])
RAYCAST = np.dtype([("P", "<f4", 3), ("dist", "<f4"), ("N", "<f4", 3), ("flags", "<u4"), ("uv", "<f4", 2),
                    ("Le", "<f4", 3), ("tri", "<u4"), ("pad", "<u4", 2)])
SPHERE = np.dtype([("center", "<f4", 3), ("radius", "<f4"), ("emission", "<f4", 3), ("flags", "<u4"),
                   ("normal_center", "<f4", 3), ("pad", "<f4")])
MISS = 0xFFFFFFFF
MODE_COMPAT, MODE_FIXED, MODE_BRUTEFORCE = 0, 1, 2
FILM_RGBAZ, FILM_SUM = 0, 1
CFG_COUNTERS = 1
RENDER_ACCUMULATE = 1
RENDER_BF_SHADOW = 2
ISECT_BVH2 = 1
STATIC_MAX_RAYS = 65536  # VMN_STATIC_MAX_RAYS
STATUS = {0: "OK", 1: "INVALID", 2: "CUDA", 3: "OOM", 4: "STATE", 5: "NCCL"}
DIST_SPLIT_SAMPLES, DIST_SPLIT_TILES = 0, 1
DIST_SCENE_BROADCAST = 1
DIST_RENDER_ASSEMBLE = 0x100
DIST_MAX_RANKS, DIST_ID_BYTES = 16, 128

EXPORTS = ["vmn_abi_version", "vmn_create", "vmn_destroy", "vmn_last_error", "vmn_set_stream", "vmn_synchronize",
           "vmn_scene_set_triangles", "vmn_scene_set_spheres", "vmn_reference_spheres", "vmn_scene_set_texture",
           "vmn_scene_commit", "vmn_intersect", "vmn_intersect_device", "vmn_intersect_static", "vmn_occluded", "vmn_raycast", "vmn_radiance",
           "vmn_camera_set", "vmn_camera_rays", "vmn_render", "vmn_read_image", "vmn_film_device",
           "vmn_film_resolve_device", "vmn_film_upload", "vmn_film_aux_device", "vmn_read_image_rgba8", "vmn_get_stats",
           "vmn_probe_read_bandwidth",
           "vmn_dist_create_local", "vmn_dist_unique_id", "vmn_dist_create_rank", "vmn_dist_destroy", "vmn_dist_last_error", "vmn_dist_world",
           "vmn_dist_rank", "vmn_dist_local_count", "vmn_dist_context", "vmn_dist_scene_set_triangles", "vmn_dist_scene_set_spheres",
           "vmn_dist_scene_set_texture", "vmn_dist_scene_commit", "vmn_dist_camera_set", "vmn_dist_render", "vmn_dist_read_image",
           "vmn_dist_get_stats"]


def kernel_source_hash():
    """sha256 (first 16 hex digits) over the CUDA sources of the library, in name order: stamps measurements that are only valid for
    the kernels they were captured on (profiles/traffic.json -> bench.py roofline.traffic)."""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(_HERE, "csrc", "*.cu")) + glob.glob(os.path.join(_HERE, "csrc", "*.cuh"))):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("flags", C.c_uint32), ("max_paths", C.c_uint64), ("stream", C.c_void_p)]


class Camera(C.Structure):
    """vmn_camera = the fields of Vermilion::Camera the integrator reads (camera.h:66-104), after its constructor."""
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("position", C.c_float * 3), ("rotation", C.c_float * 3),
                ("dist_to_film", C.c_float), ("sensor_x", C.c_float), ("sensor_y", C.c_float), ("spp", C.c_uint32)]


class RenderParams(C.Structure):
    _fields_ = [("mode", C.c_int32), ("max_segments", C.c_uint32), ("seed", C.c_uint64), ("pixel_begin", C.c_uint64),
                ("pixel_end", C.c_uint64), ("sample_begin", C.c_uint32), ("sample_end", C.c_uint32), ("tile_size", C.c_uint32),
                ("tile_rank", C.c_uint32), ("tile_world", C.c_uint32), ("flags", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_nodes", C.c_uint64), ("n_wide_nodes", C.c_uint64), ("node_bytes", C.c_uint64),
                ("tri_bytes", C.c_uint64), ("build_ms", C.c_double), ("rays_traced", C.c_uint64), ("paths_traced", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64), ("extend_ms", C.c_double), ("shade_ms", C.c_double),
                ("raygen_ms", C.c_double), ("resolve_ms", C.c_double), ("total_ms", C.c_double), ("kernel_launches", C.c_uint64),
                ("wide_depth", C.c_uint64), ("l2_persist_bytes", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class DistStats(C.Structure):
    _fields_ = [("world", C.c_uint32), ("local", C.c_uint32), ("render_ms", C.c_double * 16), ("frame_ms", C.c_double),
                ("collective_ms", C.c_double), ("scene_ms", C.c_double), ("rays_traced", C.c_uint64)]


def camera_from_settings(width, height, position, rotation_deg, back_distance=6.0, back_size=(3.6, 2.4), spp=16):
    """Vermilion::Camera's constructor (camera.cpp:34-56): degrees -> radians with pi = 3.1415926535 in double, x and y negated."""
    cam = Camera()
    cam.width, cam.height = width, height
    rot = [-float(rotation_deg[0]) * 3.1415926535 / 180, -float(rotation_deg[1]) * 3.1415926535 / 180,
           float(rotation_deg[2]) * 3.1415926535 / 180]
    for i in range(3):
        cam.position[i] = position[i]
        cam.rotation[i] = np.float32(rot[i])
    cam.dist_to_film = back_distance
    cam.sensor_x, cam.sensor_y = back_size
    cam.spp = spp
    return cam


class VmnError(RuntimeError):
    pass


_LIB = None


def load_library():
    """dlopen the C-ABI library and declare its prototypes. Loads without a GPU (no compute call is made)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise VmnError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32
    L.vmn_abi_version.restype = C.c_int
    L.vmn_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.vmn_destroy.argtypes = [vp]
    L.vmn_destroy.restype = None
    L.vmn_last_error.argtypes = [vp]
    L.vmn_last_error.restype = C.c_char_p
    L.vmn_set_stream.argtypes = [vp, vp]
    L.vmn_synchronize.argtypes = [vp]
    L.vmn_scene_set_triangles.argtypes = [vp, vp, vp, vp, u64]
    L.vmn_scene_set_spheres.argtypes = [vp, vp, u32]
    L.vmn_reference_spheres.argtypes = [vp]
    L.vmn_reference_spheres.restype = None
    L.vmn_scene_set_texture.argtypes = [vp, u32, vp, u32, u32, u32]
    L.vmn_scene_commit.argtypes = [vp]
    L.vmn_intersect.argtypes = [vp, vp, u64, vp, u32]
    L.vmn_intersect_device.argtypes = [vp, vp, u64, vp, u32]
    L.vmn_intersect_static.argtypes = [vp, vp, u64, vp, u32]
    L.vmn_occluded.argtypes = [vp, vp, u64, vp]
    L.vmn_raycast.argtypes = [vp, vp, u64, vp]
    L.vmn_radiance.argtypes = [vp, vp, u64, u64, i32, u32, vp, vp]
    L.vmn_camera_set.argtypes = [vp, C.POINTER(Camera)]
    L.vmn_camera_rays.argtypes = [vp, vp, vp, vp, vp, vp, u64, vp]
    L.vmn_render.argtypes = [vp, C.POINTER(RenderParams)]
    L.vmn_read_image.argtypes = [vp, vp, i32]
    L.vmn_film_device.argtypes = [vp, C.POINTER(vp), C.POINTER(u64)]
    L.vmn_film_resolve_device.argtypes = [vp, vp, vp]
    L.vmn_film_upload.argtypes = [vp, vp]
    L.vmn_film_aux_device.argtypes = [vp, C.POINTER(vp), C.POINTER(u64)]
    L.vmn_read_image_rgba8.argtypes = [vp, vp, vp]
    L.vmn_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.vmn_probe_read_bandwidth.argtypes = [vp, u64, u32, C.POINTER(C.c_double)]
    L.vmn_dist_create_local.argtypes = [vp, u32, u32, u64, C.POINTER(vp)]
    L.vmn_dist_unique_id.argtypes = [vp]
    L.vmn_dist_create_rank.argtypes = [vp, u32, u32, C.POINTER(Config), C.POINTER(vp)]
    L.vmn_dist_destroy.argtypes = [vp]
    L.vmn_dist_destroy.restype = None
    L.vmn_dist_last_error.argtypes = [vp]
    L.vmn_dist_last_error.restype = C.c_char_p
    for f in (L.vmn_dist_world, L.vmn_dist_rank, L.vmn_dist_local_count):
        f.argtypes = [vp]
        f.restype = u32
    L.vmn_dist_context.argtypes = [vp, u32]
    L.vmn_dist_context.restype = vp
    L.vmn_dist_scene_set_triangles.argtypes = [vp, vp, vp, vp, u64, u32]
    L.vmn_dist_scene_set_spheres.argtypes = [vp, vp, u32]
    L.vmn_dist_scene_set_texture.argtypes = [vp, u32, vp, u32, u32, u32]
    L.vmn_dist_scene_commit.argtypes = [vp]
    L.vmn_dist_camera_set.argtypes = [vp, C.POINTER(Camera)]
    L.vmn_dist_render.argtypes = [vp, C.POINTER(RenderParams), i32, u32]
    L.vmn_dist_read_image.argtypes = [vp, vp, i32]
    L.vmn_dist_get_stats.argtypes = [vp, C.POINTER(DistStats)]
    _LIB = L
    return L


def reference_spheres():
    s = np.zeros(8, SPHERE)
    load_library().vmn_reference_spheres(s.ctypes.data_as(C.c_void_p))
    return s


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Engine:
    """One vmn_ctx on one CUDA device."""

    def __init__(self, device=0, counters=False, max_paths=0, stream=None):
        self.L = load_library()
        self.h = C.c_void_p()
        cfg = Config(device, CFG_COUNTERS if counters else 0, max_paths, stream)
        st = self.L.vmn_create(C.byref(cfg), C.byref(self.h))
        if st != 0:
            msg = self.L.vmn_last_error(self.h).decode() if self.h else "allocation failed"
            if self.h:
                self.L.vmn_destroy(self.h)
                self.h = C.c_void_p()
            raise VmnError(f"vmn_create: {STATUS.get(st, st)}: {msg}")

    def _chk(self, st, what):
        if st != 0:
            raise VmnError(f"{what}: {STATUS.get(st, st)}: {self.L.vmn_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.vmn_destroy(self.h)
            self.h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # scene ------------------------------------------------------------------------------------------
    def set_texture(self, slot, texture):
        """MeshEngine::boundTextures[slot] (slot 0 or 1); texture: float32 [h, w, c] or None to unbind."""
        if texture is None:
            self._chk(self.L.vmn_scene_set_texture(self.h, slot, None, 0, 0, 0), "vmn_scene_set_texture")
            return
        tex = np.ascontiguousarray(texture, np.float32)
        h, w, c = tex.shape
        self._chk(self.L.vmn_scene_set_texture(self.h, slot, _p(tex), w, h, c), "vmn_scene_set_texture")

    def set_scene(self, pos, nrm=None, uv=None, spheres=None, texture=None, texture1=None):
        """spheres: None = the reference's eight (meshEngine.cpp:377-500); an empty array disables them."""
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 9)
        nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float32).reshape(-1, 9)
        uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 6)
        self._chk(self.L.vmn_scene_set_triangles(self.h, _p(pos), _p(nrm), _p(uv), len(pos)), "vmn_scene_set_triangles")
        sp = reference_spheres() if spheres is None else np.ascontiguousarray(spheres, SPHERE)
        self._chk(self.L.vmn_scene_set_spheres(self.h, _p(sp) if len(sp) else None, len(sp)), "vmn_scene_set_spheres")
        if texture is not None:
            self.set_texture(0, texture)
        if texture1 is not None:
            self.set_texture(1, texture1)
        self._chk(self.L.vmn_scene_commit(self.h), "vmn_scene_commit")

    # ray queries ------------------------------------------------------------------------------------
    def intersect(self, rays, flags=0):
        rays = np.ascontiguousarray(rays).view(RAY).reshape(-1)
        out = np.empty(len(rays), HIT)
        self._chk(self.L.vmn_intersect(self.h, _p(rays), len(rays), _p(out), flags), "vmn_intersect")
        return out

    def intersect_device(self, d_rays_ptr, n, d_out_ptr, flags=0):
        self._chk(self.L.vmn_intersect_device(self.h, d_rays_ptr, n, d_out_ptr, flags), "vmn_intersect_device")

    def intersect_static(self, rays, flags=0):
        """vmn_intersect_static: as intersect(), with the device scratch in the library's static array (no allocation; at most
        STATIC_MAX_RAYS rays per call)."""
        rays = np.ascontiguousarray(rays).view(RAY).reshape(-1)
        out = np.empty(len(rays), HIT)
        self._chk(self.L.vmn_intersect_static(self.h, _p(rays), len(rays), _p(out), flags), "vmn_intersect_static")
        return out

    def occluded(self, rays):
        rays = np.ascontiguousarray(rays).view(RAY).reshape(-1)
        out = np.empty(len(rays), np.uint8)
        self._chk(self.L.vmn_occluded(self.h, _p(rays), len(rays), _p(out)), "vmn_occluded")
        return out

    def raycast(self, rays):
        rays = np.ascontiguousarray(rays).view(RAY).reshape(-1)
        out = np.empty(len(rays), RAYCAST)
        self._chk(self.L.vmn_raycast(self.h, _p(rays), len(rays), _p(out)), "vmn_raycast")
        return out

    def radiance(self, rays, seed, mode=MODE_COMPAT, max_segments=5):
        rays = np.ascontiguousarray(rays).view(RAY).reshape(-1)
        out = np.empty((len(rays), 4), np.float32)
        cnt = np.empty(len(rays), np.uint32)
        self._chk(self.L.vmn_radiance(self.h, _p(rays), len(rays), seed, mode, max_segments, _p(out), _p(cnt)), "vmn_radiance")
        return out, cnt

    # camera / render --------------------------------------------------------------------------------
    def set_camera(self, cam):
        self.cam = cam
        self._chk(self.L.vmn_camera_set(self.h, C.byref(cam)), "vmn_camera_set")

    def camera_rays(self, pixel, sx, sy, jx, jy):
        pixel = np.ascontiguousarray(pixel, np.uint64)
        sx = np.ascontiguousarray(sx, np.uint32)
        sy = np.ascontiguousarray(sy, np.uint32)
        jx = np.ascontiguousarray(jx, np.float32)
        jy = np.ascontiguousarray(jy, np.float32)
        out = np.empty(len(pixel), RAY)
        self._chk(self.L.vmn_camera_rays(self.h, _p(pixel), _p(sx), _p(sy), _p(jx), _p(jy), len(pixel), _p(out)), "vmn_camera_rays")
        return out

    def render(self, mode=MODE_FIXED, seed=1, max_segments=5, pixel_begin=0, pixel_end=0, sample_begin=0, sample_end=0, tile_size=0,
               tile_rank=0, tile_world=1, accumulate=False, shadow=False):
        prm = RenderParams(mode, max_segments, seed, pixel_begin, pixel_end, sample_begin, sample_end, tile_size, tile_rank, tile_world,
                           (RENDER_ACCUMULATE if accumulate else 0) | (RENDER_BF_SHADOW if shadow else 0))
        self._chk(self.L.vmn_render(self.h, C.byref(prm)), "vmn_render")

    def read_image(self, layout=FILM_RGBAZ, out=None):
        """out: optional preallocated float32 array (e.g. pinned host memory) of shape [pixels, 5 | 4]."""
        npx = self.cam.width * self.cam.height
        shape = (npx, 5 if layout == FILM_RGBAZ else 4)
        if out is None:
            out = np.empty(shape, np.float32)
        assert out.dtype == np.float32 and out.shape == shape and out.flags.c_contiguous
        self._chk(self.L.vmn_read_image(self.h, _p(out), layout), "vmn_read_image")
        return out

    def read_image_rgba8(self, want_depth=True):
        """Camera::saveFrame's 8-bit conversion (camera.cpp:138-175) done on the device: (uint8 [pixels, 4], float32 [pixels] | None)."""
        npx = self.cam.width * self.cam.height
        rgba = np.empty((npx, 4), np.uint8)
        depth = np.empty(npx, np.float32) if want_depth else None
        self._chk(self.L.vmn_read_image_rgba8(self.h, _p(rgba), _p(depth)), "vmn_read_image_rgba8")
        return rgba, depth

    def film_device(self):
        ptr, n = C.c_void_p(), C.c_uint64()
        self._chk(self.L.vmn_film_device(self.h, C.byref(ptr), C.byref(n)), "vmn_film_device")
        return ptr.value, n.value

    def upload_film(self, film_sum):
        """restore the raw sum film (FILM_SUM layout) — resume point of a progressive render"""
        film_sum = np.ascontiguousarray(film_sum, np.float32)
        assert film_sum.shape == (self.cam.width * self.cam.height, 4)
        self._chk(self.L.vmn_film_upload(self.h, _p(film_sum)), "vmn_film_upload")

    def film_aux_device(self):
        ptr, n = C.c_void_p(), C.c_uint64()
        self._chk(self.L.vmn_film_aux_device(self.h, C.byref(ptr), C.byref(n)), "vmn_film_aux_device")
        return ptr.value, n.value

    def resolve_device(self, d_sum_ptr, out=None):
        npx = self.cam.width * self.cam.height
        if out is None:
            out = np.empty((npx, 5), np.float32)
        assert out.dtype == np.float32 and out.shape == (npx, 5) and out.flags.c_contiguous
        self._chk(self.L.vmn_film_resolve_device(self.h, d_sum_ptr, _p(out)), "vmn_film_resolve_device")
        return out

    def set_stream(self, stream_ptr):
        self._chk(self.L.vmn_set_stream(self.h, stream_ptr), "vmn_set_stream")

    def synchronize(self):
        self._chk(self.L.vmn_synchronize(self.h), "vmn_synchronize")

    def probe_read_bandwidth(self, nbytes, passes):
        """GB/s of 128-bit reads over a working set of nbytes (vmn_probe_read_bandwidth): L2 when it fits the L2, else HBM."""
        g = C.c_double()
        self._chk(self.L.vmn_probe_read_bandwidth(self.h, C.c_uint64(nbytes), C.c_uint32(passes), C.byref(g)), "vmn_probe_read_bandwidth")
        return g.value

    def stats(self):
        s = Stats()
        self._chk(self.L.vmn_get_stats(self.h, C.byref(s)), "vmn_get_stats")
        return s.as_dict()


class DistGroup:
    """One vmn_dist group (include/vermilion_b200.h): several B200s behind one render call. local(devices): one process, one member per
    device; rank(id, rank, world, ...): one process per device (torchrun), `id` = unique_id() of rank 0 handed to every rank."""

    def __init__(self, handle, lib):
        self.h, self.L, self.cam = handle, lib, None

    @staticmethod
    def unique_id():
        L = load_library()
        buf = (C.c_uint8 * DIST_ID_BYTES)()
        st = L.vmn_dist_unique_id(buf)
        if st != 0:
            raise VmnError(f"vmn_dist_unique_id: {STATUS.get(st, st)} (libnccl.so.2 not loadable?)")
        return bytes(buf)

    @classmethod
    def _made(cls, L, st, h, what):
        if st != 0:
            msg = L.vmn_dist_last_error(h).decode() if h else "allocation failed"
            if h:
                L.vmn_dist_destroy(h)
            raise VmnError(f"{what}: {STATUS.get(st, st)}: {msg}")
        return cls(h, L)

    @classmethod
    def local(cls, devices, counters=False, max_paths=0):
        L = load_library()
        h = C.c_void_p()
        arr = (C.c_int32 * len(devices))(*devices)
        st = L.vmn_dist_create_local(arr, len(devices), CFG_COUNTERS if counters else 0, max_paths, C.byref(h))
        return cls._made(L, st, h, "vmn_dist_create_local")

    @classmethod
    def rank(cls, uid, rank, world, device=0, counters=False, max_paths=0, stream=None):
        L = load_library()
        h = C.c_void_p()
        cfg = Config(device, CFG_COUNTERS if counters else 0, max_paths, stream)
        buf = (C.c_uint8 * DIST_ID_BYTES)(*uid) if uid is not None else None
        st = L.vmn_dist_create_rank(buf, rank, world, C.byref(cfg), C.byref(h))
        return cls._made(L, st, h, "vmn_dist_create_rank")

    def _chk(self, st, what):
        if st != 0:
            raise VmnError(f"{what}: {STATUS.get(st, st)}: {self.L.vmn_dist_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.vmn_dist_destroy(self.h)
            self.h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def world(self):
        return self.L.vmn_dist_world(self.h)

    @property
    def my_rank(self):
        return self.L.vmn_dist_rank(self.h)

    def set_scene(self, pos, nrm=None, uv=None, spheres=None, texture=None, texture1=None, broadcast=False):
        """pos / nrm / uv may be None on the non-root ranks of a rank group when broadcast=True (the root's arrays cross NVLink)."""
        pos = None if pos is None else np.ascontiguousarray(pos, np.float32).reshape(-1, 9)
        nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float32).reshape(-1, 9)
        uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 6)
        n = 0 if pos is None else len(pos)
        self._chk(self.L.vmn_dist_scene_set_triangles(self.h, _p(pos), _p(nrm), _p(uv), n, DIST_SCENE_BROADCAST if broadcast else 0),
                  "vmn_dist_scene_set_triangles")
        sp = reference_spheres() if spheres is None else np.ascontiguousarray(spheres, SPHERE)
        self._chk(self.L.vmn_dist_scene_set_spheres(self.h, _p(sp) if len(sp) else None, len(sp)), "vmn_dist_scene_set_spheres")
        for slot, tex in ((0, texture), (1, texture1)):
            if tex is not None:
                t = np.ascontiguousarray(tex, np.float32)
                self._chk(self.L.vmn_dist_scene_set_texture(self.h, slot, _p(t), t.shape[1], t.shape[0], t.shape[2]), "vmn_dist_scene_set_texture")
        self._chk(self.L.vmn_dist_scene_commit(self.h), "vmn_dist_scene_commit")

    def set_camera(self, cam):
        self.cam = cam
        self._chk(self.L.vmn_dist_camera_set(self.h, C.byref(cam)), "vmn_dist_camera_set")

    def render(self, mode=MODE_FIXED, seed=1, max_segments=5, split=DIST_SPLIT_SAMPLES, tile=64, sample_begin=0, sample_end=0,
               accumulate=False, assemble=False, shadow=False):
        flags = (RENDER_ACCUMULATE if accumulate else 0) | (RENDER_BF_SHADOW if shadow else 0) | (DIST_RENDER_ASSEMBLE if assemble else 0)
        prm = RenderParams(mode, max_segments, seed, 0, 0, sample_begin, sample_end, 0, 0, 0, flags)
        self._chk(self.L.vmn_dist_render(self.h, C.byref(prm), split, tile), "vmn_dist_render")

    def read_image(self, layout=FILM_RGBAZ, out=None):
        """the assembled frame on the root; None on the other ranks of a rank group"""
        if self.my_rank != 0:
            self._chk(self.L.vmn_dist_read_image(self.h, None, layout), "vmn_dist_read_image")
            return None
        npx = self.cam.width * self.cam.height
        shape = (npx, 5 if layout == FILM_RGBAZ else 4)
        if out is None:
            out = np.empty(shape, np.float32)
        assert out.dtype == np.float32 and out.shape == shape and out.flags.c_contiguous
        self._chk(self.L.vmn_dist_read_image(self.h, _p(out), layout), "vmn_dist_read_image")
        return out

    def stats(self):
        s = DistStats()
        self._chk(self.L.vmn_dist_get_stats(self.h, C.byref(s)), "vmn_dist_get_stats")
        d = {k: getattr(s, k) for k, _ in s._fields_ if k != "render_ms"}
        d["render_ms"] = list(s.render_ms)[:s.local]
        return d

    def member_stats(self, i=0):
        st = Stats()
        ctx = self.L.vmn_dist_context(self.h, i)
        if not ctx or self.L.vmn_get_stats(ctx, C.byref(st)) != 0:
            raise VmnError("vmn_get_stats on a group member")
        return st.as_dict()
