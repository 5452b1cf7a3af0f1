"""Image writer (include/minipath_b200.h: mp_image_write_png / _ppm; SURVEY §8(f).4 — the reference never saves its image):
the PNG must decode back to the same bytes with an independent decoder (zlib + the PNG chunk layout), CRCs included."""
import struct
import zlib

import numpy as np

import minipath_b200 as mp


def _decode_png(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert crc == (zlib.crc32(typ + body) & 0xFFFFFFFF), typ
        chunks.append((typ, body))
        pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    w, h, depth, ctype, comp, flt, lace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, ctype, comp, flt, lace) == (8, 6, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(h, 1 + w * 4)
    assert (raw[:, 0] == 0).all()  # filter type None on every row
    return raw[:, 1:].reshape(h, w, 4)


def test_png_and_ppm_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    for w, h in ((1, 1), (37, 5), (300, 260)):  # the last one spans several 65535-byte stored blocks
        img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        mp.save_image(tmp_path / "a.png", img)
        assert np.array_equal(_decode_png(tmp_path / "a.png"), img)
        mp.save_image(tmp_path / "a.ppm", img)
        ppm = open(tmp_path / "a.ppm", "rb").read()
        head = f"P6\n{w} {h}\n255\n".encode()
        assert ppm.startswith(head) and np.array_equal(np.frombuffer(ppm[len(head):], np.uint8).reshape(h, w, 3), img[..., :3])


def test_bad_arguments(tmp_path):
    import pytest

    with pytest.raises(mp.MpError):
        mp.save_image(tmp_path / "no_such_dir" / "a.png", np.zeros((2, 2, 4), np.uint8))
    with pytest.raises(ValueError):
        mp.save_image(tmp_path / "a.png", np.zeros((2, 2, 3), np.uint8))
