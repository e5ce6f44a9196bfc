"""Gaussian scene container, the 4-splat test scene and the 3DGS PLY loader.

Keeps the reference's data API (/root/reference/util_gau.py:6-125): ``GaussianData`` with
``flat()``, ``__len__`` and ``sh_dim``; ``naive_gaussian() -> (GaussianData, 0, 0)``;
``load_ply(path) -> (GaussianData, bounding_box, center)``.  ``plyfile`` and ``open3d`` are not
used: the PLY container is parsed here with NumPy (header on the host, rows as one structured
array), and the activations reproduce the reference's dtype path (float64 for rotation
normalisation and exp(scale), float32 for the opacity sigmoid) so results are bit-identical.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class GaussianData:
    # /root/reference/util_gau.py:6-22
    xyz: np.ndarray
    rot: np.ndarray
    scale: np.ndarray
    opacity: np.ndarray
    sh: np.ndarray

    def flat(self) -> np.ndarray:
        ret = np.concatenate([self.xyz, self.rot, self.scale, self.opacity, self.sh], axis=-1)
        return np.ascontiguousarray(ret)

    def __len__(self):
        return len(self.xyz)

    @property
    def sh_dim(self):
        return self.sh.shape[-1]


def naive_gaussian():
    """The four-splat scene of /root/reference/util_gau.py:25-60: origin + unit axes, identity
    rotations, 0.03 scales with one 0.2 axis, colours magenta/red/green/blue, opacity 1."""
    xyz = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]]).astype(np.float32)
    rot = np.tile(np.array([1, 0, 0, 0]), (4, 1)).astype(np.float32)
    s = np.full((4, 3), 0.03)
    s[1, 0] = s[2, 1] = s[3, 2] = 0.2
    s = s.astype(np.float32)
    c = np.array([[1, 0, 1], [1, 0, 0], [0, 1, 0], [0, 0, 1]]).astype(np.float32)
    c = (c - 0.5) / 0.28209
    a = np.ones((4, 1)).astype(np.float32)
    return GaussianData(xyz, rot, s, a, c), 0, 0


# ------------------------------------------------------------------------------------------------
# PLY container
# ------------------------------------------------------------------------------------------------
_PLY_TYPES = {
    "char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1",
    "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
    "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
    "float": "f4", "float32": "f4", "double": "f8", "float64": "f8",
}


def read_ply_header(f):
    """Parses a PLY header from the open binary file ``f`` (positioned at byte 0) and leaves ``f`` at
    the first body byte.  Returns (format, vertex_count, [(property name, numpy type code), ...]) for the
    FIRST element, the one load_ply reads (util_gau.py:66 ``plydata.elements[0]``)."""
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt = None
    count = None
    props = []
    in_first = False
    seen_first = False
    while True:
        line = f.readline()
        if not line:
            raise ValueError("unterminated PLY header")
        tok = line.decode("ascii", "replace").split()
        if not tok or tok[0] == "comment" or tok[0] == "obj_info":
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            if not seen_first:
                count = int(tok[2])
                in_first = True
                seen_first = True
            else:
                in_first = False
        elif tok[0] == "property" and in_first:
            if tok[1] == "list":
                raise ValueError("list properties are not supported")
            props.append((tok[2], _PLY_TYPES[tok[1]]))
        elif tok[0] == "end_header":
            break
    if fmt is None or count is None:
        raise ValueError("incomplete PLY header")
    return fmt, count, props


PLY_COLUMNS = (["x", "y", "z"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(45)] +
               ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)])


def ply_row_layout(props):
    """(row_bytes, offsets[59]) of the columns load_ply reads, in the order of ``PLY_COLUMNS`` — the
    table the device decoder (gsr_load_ply_rows) takes.  Columns are found by NAME and the numbered
    families are ordered by their numeric suffix, like util_gau.py:92-112.  Raises AssertionError for a
    file without exactly 45 ``f_rest_*`` (:94) and ValueError for a column that is not ``property float``."""
    names = [n for n, _ in props]
    offs = {}
    o = 0
    for n, t in props:
        offs[n] = (o, t)
        o += np.dtype(t).itemsize
    row_bytes = o
    extra = _numbered(names, "f_rest_")
    assert len(extra) == 45
    scale = _numbered(names, "scale_")
    rot = _numbered(names, "rot")
    if len(scale) != 3 or len(rot) != 4:
        raise ValueError("expected 3 scale_* and 4 rot* properties")
    cols = ["x", "y", "z", "f_dc_0", "f_dc_1", "f_dc_2"] + extra + ["opacity"] + scale + rot
    out = []
    for c in cols:
        if c not in offs:
            raise ValueError(f"PLY property {c} missing")
        off, t = offs[c]
        if t != "f4":
            raise ValueError(f"PLY property {c} is not `property float`; use the host load_ply")
        out.append(off)
    return row_bytes, out


def read_ply_vertices(path) -> np.ndarray:
    """Returns the first element of a PLY file as a NumPy structured array (one field per
    property).  Handles binary_little_endian, binary_big_endian and ascii; list properties are
    not supported (3DGS point clouds have none)."""
    with open(path, "rb") as f:
        try:
            fmt, count, props = read_ply_header(f)
        except ValueError as e:
            raise ValueError(f"{path}: {e}") from e
        if fmt == "ascii":
            raw = np.loadtxt(f, dtype=np.float64, max_rows=count, ndmin=2)
            out = np.empty(count, dtype=[(n, "<" + t) for n, t in props])
            for i, (n, _) in enumerate(props):
                out[n] = raw[:, i]
            return out
        end = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, end + t) for n, t in props])
        data = np.fromfile(f, dtype=dt, count=count)
        if len(data) != count:
            raise ValueError(f"{path}: truncated PLY body ({len(data)} of {count} vertices)")
        return data


def _numbered(names, prefix):
    sel = [n for n in names if n.startswith(prefix)]
    return sorted(sel, key=lambda x: int(x.split("_")[-1]))


def load_ply(path):
    """3DGS point cloud -> (GaussianData, bounding_box(2,3), center(3)).

    Follows /root/reference/util_gau.py:63-125 property by property: columns are looked up by
    NAME; the 45 ``f_rest_*`` columns are channel-major in the file and become
    coefficient-major / channel-minor (:99-100, :122-123); rotation is L2-normalised and scale
    exponentiated in float64 then rounded to float32 (:116-119); opacity goes through a float32
    sigmoid (:120-121).  A file without exactly 45 ``f_rest_*`` raises AssertionError (:94)."""
    max_sh_degree = 3
    v = read_ply_vertices(path)
    names = list(v.dtype.names)
    xyz = np.stack((np.asarray(v["x"]), np.asarray(v["y"]), np.asarray(v["z"])), axis=1)
    opacities = np.asarray(v["opacity"])[..., np.newaxis]

    bounding_box = np.array([xyz.min(axis=0), xyz.max(axis=0)])
    center = xyz.mean(axis=0)

    features_dc = np.stack([np.asarray(v["f_dc_0"]), np.asarray(v["f_dc_1"]), np.asarray(v["f_dc_2"])],
                           axis=1).astype(np.float64)
    extra_f_names = _numbered(names, "f_rest_")
    assert len(extra_f_names) == 3 * (max_sh_degree + 1) ** 2 - 3
    features_extra = np.stack([np.asarray(v[n]) for n in extra_f_names], axis=1).astype(np.float64)
    features_extra = features_extra.reshape((features_extra.shape[0], 3, (max_sh_degree + 1) ** 2 - 1))
    features_extra = np.transpose(features_extra, [0, 2, 1])

    scales = np.stack([np.asarray(v[n]) for n in _numbered(names, "scale_")], axis=1).astype(np.float64)
    rots = np.stack([np.asarray(v[n]) for n in _numbered(names, "rot")], axis=1).astype(np.float64)

    xyz = xyz.astype(np.float32)
    rots = rots / np.linalg.norm(rots, axis=-1, keepdims=True)
    rots = rots.astype(np.float32)
    scales = np.exp(scales).astype(np.float32)
    opacities = 1 / (1 + np.exp(-opacities))
    opacities = opacities.astype(np.float32)
    shs = np.concatenate([features_dc.reshape(-1, 3), features_extra.reshape(len(features_dc), -1)],
                         axis=-1).astype(np.float32)
    return GaussianData(xyz, rots, scales, opacities, shs), bounding_box, center


def save_ply(path, xyz, f_dc, f_rest, opacity_logit, log_scale, rot_raw, normals=None):
    """Writes the official 3DGS point_cloud.ply layout (62 float properties: x y z nx ny nz
    f_dc_0..2 f_rest_0..44 opacity scale_0..2 rot_0..3), PRE-activation values, binary little
    endian.  ``f_rest`` is (N, 45) channel-major as stored by the 3DGS trainer.  Used to make
    fixtures for ``load_ply``; the reference has no writer."""
    n = len(xyz)
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + \
            [f"f_rest_{i}" for i in range(f_rest.shape[1])] + ["opacity"] + \
            [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    if normals is None:
        normals = np.zeros((n, 3), dtype=np.float32)
    cols = np.concatenate([xyz, normals, f_dc, f_rest, opacity_logit.reshape(n, 1), log_scale, rot_raw],
                          axis=1).astype("<f4")
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n
    header += "".join(f"property float {nm}\n" for nm in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(cols).tobytes())
