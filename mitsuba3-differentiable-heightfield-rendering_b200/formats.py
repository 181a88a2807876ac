"""Scene-format adapters for the `heightfield` plugin: height data in and out of the single-channel
image formats Mitsuba's Bitmap reads without external libraries, and the XML / dict snippet that declares the shape in a
Mitsuba scene. Host-side file plumbing only -- nothing here is on the timed path.

Reference pointers: gridded data is ingested as a (H, W, 1) float tensor, row index = y (src/textures/bitmap.cpp:124-143,
259-263); PFM is one of Bitmap's native formats (src/core/bitmap.cpp, FileFormat::PFM: rows stored bottom-to-top, a negative
scale marks little-endian); scene dicts / XML follow src/python/python/xml.py and docs "scene format".
"""
from __future__ import annotations

import os
import re

import numpy as np

__all__ = ["load_heights", "save_heights", "mitsuba_dict", "mitsuba_xml"]


def _read_tokens(f, n):
    """n whitespace-separated header tokens of a netpbm / PFM file, '#' comments skipped."""
    out, tok = [], b""
    while len(out) < n:
        c = f.read(1)
        if not c:
            raise ValueError("truncated header")
        if c == b"#":
            f.readline()
            c = b" "
        if c.isspace():
            if tok:
                out.append(tok)
                tok = b""
        else:
            tok += c
    return out


def load_heights(path: str) -> np.ndarray:
    """(H, W) float32, row 0 = first image row (y = -1 edge of the object-space square).
    .npy: any real dtype, (H, W) or (H, W, 1). .pfm: 'Pf' single channel. .pgm: P5 (8 / 16 bit, scaled to [0, 1]) or P2."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".npy":
        a = np.load(path)
    elif ext == ".pfm":
        with open(path, "rb") as f:
            magic, w, h, scale = _read_tokens(f, 4)
            if magic != b"Pf":
                raise ValueError("heightfield: PFM must be single-channel ('Pf')")
            w, h, scale = int(w), int(h), float(scale)
            data = np.frombuffer(f.read(w * h * 4), dtype="<f4" if scale < 0 else ">f4")
            if data.size != w * h:
                raise ValueError("truncated PFM")
            a = data.reshape(h, w)[::-1]                      # PFM stores the bottom row first
    elif ext == ".pgm":
        with open(path, "rb") as f:
            magic, w, h, maxv = _read_tokens(f, 4)
            w, h, maxv = int(w), int(h), int(maxv)
            if magic == b"P5":
                dt = ">u2" if maxv > 255 else "u1"
                data = np.frombuffer(f.read(w * h * np.dtype(dt).itemsize), dtype=dt)
            elif magic == b"P2":
                data = np.array(f.read().split(), dtype=np.float64)
            else:
                raise ValueError("heightfield: PGM must be P5 or P2")
            if data.size != w * h:
                raise ValueError("truncated PGM")
            a = data.reshape(h, w).astype(np.float64) / maxv
    else:
        raise ValueError(f"heightfield: unsupported height file '{path}' (.npy, .pfm, .pgm)")
    a = np.asarray(a)
    if a.ndim == 3 and a.shape[2] == 1:
        a = a[:, :, 0]
    if a.ndim != 2 or a.shape[0] < 2 or a.shape[1] < 2:
        raise ValueError("heightfield: height data must be (H, W) with H, W >= 2")
    return np.ascontiguousarray(a, dtype=np.float32)


def save_heights(path: str, heights, bits: int = 16) -> None:
    """Inverse of load_heights. .npy / .pfm are lossless; .pgm quantises [0, 1] to `bits` (8 or 16)."""
    a = np.asarray(heights.detach().cpu() if hasattr(heights, "detach") else heights, dtype=np.float32)
    if a.ndim != 2:
        raise ValueError("heights must be (H, W)")
    ext = os.path.splitext(path)[1].lower()
    if ext == ".npy":
        np.save(path, a)
    elif ext == ".pfm":
        with open(path, "wb") as f:
            f.write(b"Pf\n%d %d\n-1.0\n" % (a.shape[1], a.shape[0]))
            f.write(np.ascontiguousarray(a[::-1], dtype="<f4").tobytes())
    elif ext == ".pgm":
        maxv = 65535 if bits == 16 else 255
        q = np.clip(np.rint(a.astype(np.float64) * maxv), 0, maxv)
        with open(path, "wb") as f:
            f.write(b"P5\n%d %d\n%d\n" % (a.shape[1], a.shape[0], maxv))
            f.write(q.astype(">u2" if bits == 16 else "u1").tobytes())
    else:
        raise ValueError(f"unsupported height file '{path}'")


def mitsuba_dict(filename: str, max_height: float = 1.0, to_world=None, flip_normals: bool = False, bsdf: dict | None = None) -> dict:
    """The dict `mi.load_dict` takes for this shape (what INTEGRATION.md's plugin constructor consumes)."""
    d = {"type": "heightfield", "filename": filename, "max_height": float(max_height)}
    if flip_normals:
        d["flip_normals"] = True
    if to_world is not None:
        d["to_world"] = np.asarray(to_world, dtype=np.float32).reshape(-1)[:12].reshape(3, 4).tolist()
    d["bsdf"] = bsdf or {"type": "diffuse"}
    return d


def mitsuba_xml(filename: str, max_height: float = 1.0, to_world=None, flip_normals: bool = False, shape_id: str = "heightfield") -> str:
    """XML snippet declaring the shape (scene format version 3.0.0)."""
    lines = [f'<shape type="heightfield" id="{shape_id}">', f'    <string name="filename" value="{filename}"/>',
             f'    <float name="max_height" value="{float(max_height):.9g}"/>']
    if flip_normals:
        lines.append('    <boolean name="flip_normals" value="true"/>')
    if to_world is not None:
        m = np.eye(4, dtype=np.float64)
        m[:3, :] = np.asarray(to_world, dtype=np.float64).reshape(-1)[:12].reshape(3, 4)
        vals = " ".join(f"{v:.9g}" for v in m.reshape(-1))
        lines += ['    <transform name="to_world">', f'        <matrix value="{vals}"/>', '    </transform>']
    lines += ['    <bsdf type="diffuse"/>', '</shape>']
    return "\n".join(lines) + "\n"


def parse_mitsuba_xml(snippet: str) -> dict:
    """Read back what mitsuba_xml wrote (enough of the scene format for a round trip; not a general XML loader)."""
    out = {"type": "heightfield"}
    m = re.search(r'<string name="filename" value="([^"]*)"', snippet)
    if m:
        out["filename"] = m.group(1)
    m = re.search(r'<float name="max_height" value="([^"]*)"', snippet)
    if m:
        out["max_height"] = float(m.group(1))
    if re.search(r'<boolean name="flip_normals" value="true"', snippet):
        out["flip_normals"] = True
    m = re.search(r'<matrix value="([^"]*)"', snippet)
    if m:
        out["to_world"] = np.array([float(x) for x in m.group(1).split()], np.float32).reshape(4, 4)[:3].copy()
    return out
