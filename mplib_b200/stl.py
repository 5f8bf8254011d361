"""STL reader (ASCII + binary) producing (unique vertices, triangles).

Restates what the reference gets out of assimp for collision meshes
(``src/urdf_utils.cpp:82-133, 248-283``): positions are single precision, normals are dropped
(``aiComponent_NORMALS`` in the remove-component set) and identical positions are joined
(``aiProcess_JoinIdenticalVertices``), the root node's Y-up transform is not applied, and the
per-axis URDF ``scale`` multiplies each coordinate after the cast to the scalar type.
Vertices are numbered in order of first appearance.
"""
from __future__ import annotations

import struct

import numpy as np


def _parse_ascii(data: bytes):
    toks = data.split()
    verts = []
    i = 0
    n = len(toks)
    while i < n:
        if toks[i] == b"vertex":
            verts.append((float(toks[i + 1]), float(toks[i + 2]), float(toks[i + 3])))
            i += 4
        else:
            i += 1
    if len(verts) % 3 != 0 or not verts:
        raise ValueError("malformed ASCII STL")
    return np.asarray(verts, dtype=np.float32).reshape(-1, 3, 3)


def _parse_binary(data: bytes):
    if len(data) < 84:
        raise ValueError("malformed binary STL")
    (ntri,) = struct.unpack_from("<I", data, 80)
    if len(data) < 84 + 50 * ntri:
        raise ValueError("truncated binary STL")
    rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=ntri, offset=84)
    return np.array(rec["v"], dtype=np.float32)


def _is_binary(data: bytes):
    if len(data) >= 84:
        (ntri,) = struct.unpack_from("<I", data, 80)
        if len(data) == 84 + 50 * ntri:
            return True
    head = data[:512].lstrip()
    return not head.startswith(b"solid")


def load_stl(path: str, scale=(1.0, 1.0, 1.0)):
    """-> (vertices float64 [V,3], triangles int32 [T,3]); raises ValueError like the reference's
    ``std::invalid_argument`` when the file cannot be read."""
    try:
        with open(path, "rb") as f:
            data = f.read()
    except OSError as e:
        raise ValueError(f"Could not load resource {path}\nHint: the mesh directory may be wrong.") from e
    tri = _parse_binary(data) if _is_binary(data) else _parse_ascii(data)
    flat = tri.reshape(-1, 3)
    # join identical positions, numbering by first appearance
    _, first, inv = np.unique(flat.view([("", np.float32)] * 3).reshape(-1), return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")
    rank = np.empty_like(order)
    rank[order] = np.arange(order.size)
    verts32 = flat[first[order]]
    idx = rank[inv].reshape(-1, 3).astype(np.int32)
    verts = verts32.astype(np.float64) * np.asarray(scale, dtype=np.float64)[None, :]
    return verts, idx
