"""Scene-format adapters: height files round-trip, the PFM row order and endianness follow the format
Mitsuba's Bitmap reads, and the XML / dict snippets carry the plugin's properties."""
import numpy as np
import pytest

import hfb200
from hfb200 import formats as F


def test_height_files_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    h = rng.uniform(0, 1, (7, 5)).astype(np.float32)
    for ext in (".npy", ".pfm"):
        p = str(tmp_path / ("h" + ext))
        F.save_heights(p, h)
        assert np.array_equal(F.load_heights(p), h)
    for bits, tol in ((16, 0.5 / 65535), (8, 0.5 / 255)):
        p = str(tmp_path / f"h{bits}.pgm")
        F.save_heights(p, h, bits=bits)
        assert np.abs(F.load_heights(p) - h).max() <= tol * 1.0001
    np.save(str(tmp_path / "h3.npy"), h[:, :, None].astype(np.float64))          # (H, W, 1) tensors as in bitmap.cpp:259-263
    assert np.array_equal(F.load_heights(str(tmp_path / "h3.npy")), h)


def test_pfm_layout_is_bottom_up_and_endianness_follows_the_scale_sign(tmp_path):
    h = np.arange(6, dtype=np.float32).reshape(3, 2)
    p = str(tmp_path / "a.pfm")
    F.save_heights(p, h)
    raw = open(p, "rb").read()
    assert raw.startswith(b"Pf\n2 3\n-1.0\n")
    body = np.frombuffer(raw[len(b"Pf\n2 3\n-1.0\n"):], "<f4").reshape(3, 2)
    assert np.array_equal(body[0], h[2]) and np.array_equal(body[2], h[0])          # last image row first
    with open(str(tmp_path / "be.pfm"), "wb") as f:                                  # big-endian file, comment in the header
        f.write(b"Pf\n# made by hand\n2 3\n1.0\n" + h[::-1].astype(">f4").tobytes())
    assert np.array_equal(F.load_heights(str(tmp_path / "be.pfm")), h)
    with open(str(tmp_path / "a.pgm"), "wb") as f:
        f.write(b"P2\n2 2\n10\n0 5\n10 2\n")
    assert np.allclose(F.load_heights(str(tmp_path / "a.pgm")), [[0, 0.5], [1, 0.2]])


def test_bad_files_are_rejected(tmp_path):
    with open(str(tmp_path / "rgb.pfm"), "wb") as f:
        f.write(b"PF\n2 2\n-1.0\n" + bytes(48))
    with pytest.raises(ValueError, match="single-channel"):
        F.load_heights(str(tmp_path / "rgb.pfm"))
    with open(str(tmp_path / "short.pfm"), "wb") as f:
        f.write(b"Pf\n4 4\n-1.0\n" + bytes(8))
    with pytest.raises(ValueError, match="truncated"):
        F.load_heights(str(tmp_path / "short.pfm"))
    np.save(str(tmp_path / "one.npy"), np.zeros((1, 5), np.float32))
    with pytest.raises(ValueError, match=">= 2"):
        F.load_heights(str(tmp_path / "one.npy"))
    with pytest.raises(ValueError, match="unsupported"):
        F.load_heights(str(tmp_path / "h.exr"))


def test_scene_snippets_round_trip():
    tw = np.array([[0.9, 0.1, 0.0, 0.05], [-0.1, 0.9, 0.05, -0.02], [0.0, 0.02, 1.1, 0.01]], np.float32)
    xml = F.mitsuba_xml("terrain.pfm", max_height=0.3, to_world=tw, flip_normals=True)
    assert '<shape type="heightfield"' in xml and 'name="filename" value="terrain.pfm"' in xml
    back = F.parse_mitsuba_xml(xml)
    assert back["filename"] == "terrain.pfm" and back["flip_normals"] is True
    assert np.float32(back["max_height"]) == np.float32(0.3) and np.array_equal(back["to_world"], tw)
    d = F.mitsuba_dict("terrain.pfm", 0.3, tw)
    assert d["type"] == "heightfield" and d["bsdf"] == {"type": "diffuse"} and np.allclose(d["to_world"], tw)


@pytest.mark.gpu
def test_plugin_loads_from_a_pfm_file_named_in_a_scene_dict(tmp_path):
    import torch
    rng = np.random.default_rng(1)
    h = rng.uniform(0, 1, (33, 20)).astype(np.float32)
    p = str(tmp_path / "t.pfm")
    F.save_heights(p, h)
    shape = hfb200.load_dict(F.mitsuba_dict(p, 0.4))
    assert torch.equal(shape.heights.cpu(), torch.from_numpy(h)) and shape.max_height == pytest.approx(0.4)
    shape2 = hfb200.load_dict(F.parse_mitsuba_xml(F.mitsuba_xml(p, 0.4)))
    assert np.array_equal(shape2.bbox(), shape.bbox())
