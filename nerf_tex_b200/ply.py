"""PLY meshes for the instancer (the reference reads them with igl::readPLY, instancer/src/instancer.cpp:236,278,401):
vertex positions, optional per-vertex normals (nx ny nz) and texture coordinates (s t / u v / texture_u texture_v),
triangle faces (polygons are fanned).  ascii and binary_little_endian / binary_big_endian."""

from __future__ import annotations

from typing import Dict, Optional

import numpy as np

_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
          "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
          "double": "f8", "float64": "f8"}


def read_ply(path: str) -> Dict[str, Optional[np.ndarray]]:
    """{'V': [n,3] f32, 'F': [m,3] i32 (or [0,3]), 'N': [n,3] f32 or None, 'UV': [n,2] f32 or None}."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.find(b"end_header")
    if not data.startswith(b"ply") or end < 0:
        raise ValueError(f"{path}: not a PLY file")
    header = data[:end].decode("ascii", "replace").splitlines()
    body = data[data.index(b"\n", end) + 1:]
    fmt, elements = None, []
    for line in header:
        tok = line.split()
        if not tok:
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            elements.append({"name": tok[1], "count": int(tok[2]), "props": []})
        elif tok[0] == "property":
            if tok[1] == "list":
                elements[-1]["props"].append((tok[4], "list", _TYPES[tok[2]], _TYPES[tok[3]]))
            else:
                elements[-1]["props"].append((tok[2], _TYPES[tok[1]]))
    if fmt not in ("ascii", "binary_little_endian", "binary_big_endian"):
        raise ValueError(f"{path}: unsupported PLY format {fmt}")
    end_c = "<" if fmt != "binary_big_endian" else ">"
    vert, faces = None, []
    if fmt == "ascii":
        tokens = body.split()
        pos = 0
    else:
        pos = 0
    for el in elements:
        scalar = all(len(p) == 2 for p in el["props"])
        if fmt == "ascii":
            if scalar:
                k = len(el["props"])
                arr = np.array(tokens[pos:pos + k * el["count"]], dtype=np.float64).reshape(el["count"], k)
                pos += k * el["count"]
                rec = {p[0]: arr[:, i] for i, p in enumerate(el["props"])}
            else:
                rec = None
                for _ in range(el["count"]):
                    for p in el["props"]:
                        if len(p) == 2:
                            pos += 1
                        else:
                            cnt = int(tokens[pos])
                            idx = [int(x) for x in tokens[pos + 1:pos + 1 + cnt]]
                            pos += 1 + cnt
                            if el["name"] == "face":
                                faces += [[idx[0], idx[i], idx[i + 1]] for i in range(1, cnt - 1)]
        else:
            if scalar:
                dt = np.dtype([(p[0], end_c + p[1]) for p in el["props"]])
                arr = np.frombuffer(body, dtype=dt, count=el["count"], offset=pos)
                pos += dt.itemsize * el["count"]
                rec = {name: arr[name].astype(np.float64) for name in arr.dtype.names}
            else:
                rec = None
                for _ in range(el["count"]):
                    for p in el["props"]:
                        if len(p) == 2:
                            pos += np.dtype(p[1]).itemsize
                        else:
                            cnt = int(np.frombuffer(body, dtype=end_c + p[2], count=1, offset=pos)[0])
                            pos += np.dtype(p[2]).itemsize
                            idx = np.frombuffer(body, dtype=end_c + p[3], count=cnt, offset=pos).astype(np.int64)
                            pos += np.dtype(p[3]).itemsize * cnt
                            if el["name"] == "face":
                                faces += [[int(idx[0]), int(idx[i]), int(idx[i + 1])] for i in range(1, cnt - 1)]
        if el["name"] == "vertex":
            vert = rec
    if vert is None or not all(k in vert for k in "xyz"):
        raise ValueError(f"{path}: no vertex positions")
    col = lambda *names: np.stack([vert[n] for n in names], -1).astype(np.float32)
    N = col("nx", "ny", "nz") if all(k in vert for k in ("nx", "ny", "nz")) else None
    UV = None
    for a, b in (("s", "t"), ("u", "v"), ("texture_u", "texture_v")):
        if a in vert and b in vert:
            UV = col(a, b)
            break
    return {"V": col("x", "y", "z"), "F": np.asarray(faces, np.int32).reshape(-1, 3), "N": N, "UV": UV}


def write_ply(path: str, V, F=None, N=None, UV=None, binary: bool = True) -> None:
    """Writer for procedural meshes and anchor point sets (tests, tools)."""
    V = np.asarray(V, np.float32)
    cols, names = [V], ["x", "y", "z"]
    if N is not None:
        cols.append(np.asarray(N, np.float32)); names += ["nx", "ny", "nz"]
    if UV is not None:
        cols.append(np.asarray(UV, np.float32)); names += ["s", "t"]
    table = np.ascontiguousarray(np.concatenate(cols, 1), dtype="<f4")
    points_only = F is None          # an anchor point set has no face element at all (libigl's reader, which the
    F = np.zeros((0, 3), np.int32) if F is None else np.asarray(F, np.int32)      # reference uses, divides by the face count)
    head = ["ply", "format binary_little_endian 1.0" if binary else "format ascii 1.0", f"element vertex {V.shape[0]}"]
    head += [f"property float {n}" for n in names]
    if not points_only:
        head += [f"element face {F.shape[0]}", "property list uchar int vertex_indices"]
    head += ["end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(head) + "\n").encode("ascii"))
        if binary:
            f.write(table.tobytes())
            rec = np.zeros(F.shape[0], dtype=[("n", "u1"), ("i", "<i4", 3)])
            rec["n"], rec["i"] = 3, F
            f.write(rec.tobytes())
        else:
            for row in table:
                f.write((" ".join(repr(float(x)) for x in row) + "\n").encode("ascii"))
            for tri in F:
                f.write(("3 " + " ".join(str(int(i)) for i in tri) + "\n").encode("ascii"))
