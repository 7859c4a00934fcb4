"""Film output and checkpoints. CPU: the PNG / PFM writers and the checkpoint file of vermilion_b200/host/film_io.cpp
(PNG decoded back with zlib; Camera::saveFrame's own writer, OpenImageIO, is not in this image). GPU: a progressive render interrupted,
checkpointed, restored into a NEW context and resumed reproduces the uninterrupted film bit for bit."""
import struct
import zlib

import numpy as np
import pytest

from vermilion_b200 import film_io


def _decode_png(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        assert crc == zlib.crc32(typ + data)
        chunks.append((typ, data))
        pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    w, h, depth, ctype, comp, flt, inter = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, ctype, comp, flt, inter) == (8, 6, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(h, 4 * w + 1)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 4)


def test_png_and_pfm_writers(tmp_path):
    rng = np.random.default_rng(1)
    w, h = 37, 21  # ragged sizes
    rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    depth = rng.random((h, w), dtype=np.float32) * 1000
    film_io.save_frame(str(tmp_path / "frame"), rgba, depth, w, h)
    assert np.array_equal(_decode_png(tmp_path / "frame.png"), rgba)
    b = open(tmp_path / "frame_depth.pfm", "rb").read()
    head = f"Pf\n{w} {h}\n-1.0\n".encode()
    assert b.startswith(head)
    got = np.frombuffer(b[len(head):], "<f4").reshape(h, w)[::-1]  # PFM rows run bottom to top
    assert np.array_equal(got, depth)
    film_io.write_pfm(str(tmp_path / "rgb.pfm"), rng.random((h, w, 3), dtype=np.float32), w, h, 3)
    assert open(tmp_path / "rgb.pfm", "rb").read().startswith(b"PF\n")
    with pytest.raises(OSError):
        film_io.save_frame(str(tmp_path / "no_such_dir" / "x"), rgba, depth, w, h)


def test_checkpoint_round_trip_and_corruption(tmp_path):
    rng = np.random.default_rng(2)
    w, h = 19, 11
    film = rng.random((w * h, 4), dtype=np.float32)
    p = str(tmp_path / "film.vmn")
    film_io.checkpoint_save(p, film, w, h, spp_done=24, spp_total=64, seed=77, max_segments=5)
    got, hdr = film_io.checkpoint_load(p)
    assert np.array_equal(got, film)
    assert (hdr.width, hdr.height, hdr.spp_done, hdr.spp_total, hdr.seed, hdr.max_segments) == (w, h, 24, 64, 77, 5)
    assert not (tmp_path / "film.vmn.tmp").exists()  # written to a temporary and renamed
    raw = bytearray(open(p, "rb").read())
    raw[-5] ^= 0x40  # flip a payload bit: CRC mismatch
    open(p, "wb").write(raw)
    with pytest.raises(OSError):
        film_io.checkpoint_load(p)
    open(p, "wb").write(raw[:-8])  # truncated
    assert film_io.checkpoint_info(p) is None
    assert film_io.checkpoint_info(str(tmp_path / "absent")) is None


@pytest.mark.gpu
def test_checkpoint_resume_reproduces_the_uninterrupted_film(tmp_path):
    from vermilion_b200 import engine, scenes
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    cam = engine.camera_from_settings(64, 48, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.7), 12)
    p = str(tmp_path / "film.vmn")
    with engine.Engine(device=0) as eng:
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(cam)
        eng.render(mode=engine.MODE_FIXED, seed=5)
        full = eng.read_image(engine.FILM_SUM)
        full_img = eng.read_image()
        eng.render(mode=engine.MODE_FIXED, seed=5, sample_begin=0, sample_end=5)  # "interrupted" after 5 of 12 spp
        film_io.checkpoint_save(p, eng.read_image(engine.FILM_SUM), cam.width, cam.height, 5, cam.spp, 5)
    film, hdr = film_io.checkpoint_load(p)
    with engine.Engine(device=0) as eng:  # a new context, as after a restart
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(cam)
        eng.upload_film(film)
        eng.render(mode=engine.MODE_FIXED, seed=hdr.seed, sample_begin=hdr.spp_done, sample_end=hdr.spp_total, accumulate=True)
        assert np.array_equal(eng.read_image(engine.FILM_SUM), full)
        rgba, depth = eng.read_image_rgba8()
        film_io.save_frame(str(tmp_path / "out"), rgba, depth, cam.width, cam.height)
    assert np.array_equal(_decode_png(tmp_path / "out.png").reshape(-1, 4), np.floor(full_img[:, :4] * np.float32(255)).astype(np.uint8))
