"""ctypes binding of vermilion_b200/host/film_io.h (libvermilion_host.so): PNG / PFM output of the film and progressive-render
checkpoints. Host code only; see the header for what each call replaces in the reference (Camera::saveFrame, camera.cpp:126-194)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class CheckpointHeader(C.Structure):
    _fields_ = [("magic", C.c_char * 8), ("width", C.c_uint32), ("height", C.c_uint32), ("spp_done", C.c_uint32), ("spp_total", C.c_uint32),
                ("seed", C.c_uint64), ("max_segments", C.c_uint32), ("crc32", C.c_uint32), ("frame_fingerprint", C.c_uint64)]


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(os.path.join(_HERE, "host", "libvermilion_host.so"))
        L.vmnh_write_png_rgba8.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.vmnh_write_pfm.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        L.vmnh_save_frame.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.vmnh_checkpoint_save.argtypes = [C.c_char_p, C.c_void_p, C.POINTER(CheckpointHeader)]
        L.vmnh_checkpoint_info.argtypes = [C.c_char_p, C.POINTER(CheckpointHeader)]
        L.vmnh_checkpoint_load.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.POINTER(CheckpointHeader)]
        _LIB = L
    return _LIB


def _chk(rc, what):
    if rc != 0:
        raise OSError(f"{what} failed with code {rc}")


def save_frame(name, rgba8, depth, width, height):
    """Camera::saveFrame for an RGBAZ film: <name>.png (8-bit RGBA) and <name>_depth.pfm"""
    rgba8 = np.ascontiguousarray(rgba8, np.uint8)
    depth = None if depth is None else np.ascontiguousarray(depth, np.float32)
    assert rgba8.size == width * height * 4
    _chk(lib().vmnh_save_frame(os.fsencode(name), rgba8.ctypes.data, None if depth is None else depth.ctypes.data, width, height), "vmnh_save_frame")


def write_pfm(path, data, width, height, channels):
    data = np.ascontiguousarray(data, np.float32)
    _chk(lib().vmnh_write_pfm(os.fsencode(path), data.ctypes.data, width, height, channels), "vmnh_write_pfm")


def checkpoint_save(path, film_sum, width, height, spp_done, spp_total, seed, max_segments=5, frame_fingerprint=0):
    film_sum = np.ascontiguousarray(film_sum, np.float32)
    assert film_sum.size == width * height * 4
    h = CheckpointHeader(b"VMNFILM2", width, height, spp_done, spp_total, seed, max_segments, 0, frame_fingerprint)
    _chk(lib().vmnh_checkpoint_save(os.fsencode(path), film_sum.ctypes.data, C.byref(h)), "vmnh_checkpoint_save")


def checkpoint_info(path):
    h = CheckpointHeader()
    rc = lib().vmnh_checkpoint_info(os.fsencode(path), C.byref(h))
    return None if rc != 0 else h


def checkpoint_load(path):
    h = checkpoint_info(path)
    if h is None:
        raise OSError(f"{path}: not a film checkpoint")
    film = np.empty((h.width * h.height, 4), np.float32)
    _chk(lib().vmnh_checkpoint_load(os.fsencode(path), film.ctypes.data, film.size, C.byref(h)), "vmnh_checkpoint_load")
    return film, h
