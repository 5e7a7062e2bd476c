"""On-disk interchange with the reference: the Gaussian PLY of scene/io.py:13-132.

Vertex properties, all float32, in the reference's order: x y z opacity scale_0..2 rot_0..3 then either
f_dc_0..2 f_rest_0..(3K-4) (with_sh: SH coefficients (N,3,K), DC first, rest channel-major) or f_0..f_{D-1} (a plain
(N,D) feature plane, e.g. the MLP scene's point features).  opacity = alpha_logit, scale = log_scaling, rot = the
NORMALISED quaternion in wxyz order (the scene keeps xyzw).  The file is `binary_little_endian 1.0`, which is what
plyfile (absent here) writes by default for PlyData([element]); the byte layout below follows the PLY specification."""
import numpy as np
import torch

_BASE = ["x", "y", "z", "opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1", "char": "i1",
              "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2", "int": "<i4", "int32": "<i4",
              "uint": "<u4", "uint32": "<u4"}


def _np(t):
  return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def to_vertex_array(gaussians, with_sh=True):
  """reference to_plydata (scene/io.py:13-68) -> the structured vertex array.  gaussians: dict position (N,3),
  log_scaling (N,3), rotation (N,4) xyzw, alpha_logit (N,) or (N,1), and sh (N,3,K) [with_sh] or feature (N,D)."""
  pos, ls = _np(gaussians["position"]).astype(np.float32), _np(gaussians["log_scaling"]).astype(np.float32)
  rot = _np(gaussians["rotation"]).astype(np.float32)
  alpha = _np(gaussians["alpha_logit"]).astype(np.float32).reshape(-1)
  n = pos.shape[0]
  if with_sh:
    sh = _np(gaussians["sh"] if "sh" in gaussians else gaussians["feature"]).astype(np.float32)
    assert sh.ndim == 3, f"expected (N,3,K) SH coefficients, got {sh.shape}"
    names = ["f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(sh.shape[1] * sh.shape[2] - 3)]
    cols = np.concatenate([sh[:, :, 0].reshape(n, 3), sh[:, :, 1:].reshape(n, -1)], axis=1)
  else:
    feat = _np(gaussians["feature"]).astype(np.float32)
    assert feat.ndim == 2, f"expected an (N,D) feature plane, got {feat.shape}"
    names, cols = [f"f_{i}" for i in range(feat.shape[1])], feat
  vertex = np.zeros(n, dtype=[(k, "<f4") for k in _BASE + names])
  for i, k in enumerate("xyz"):
    vertex[k] = pos[:, i]
  vertex["opacity"] = alpha
  # F.normalize (eps 1e-12) then xyzw -> wxyz
  q = rot / np.maximum(np.sqrt((rot * rot).sum(axis=1, keepdims=True, dtype=np.float32)), np.float32(1e-12))
  q = np.roll(q, 1, axis=1)
  for i in range(3):
    vertex[f"scale_{i}"] = ls[:, i]
  for i in range(4):
    vertex[f"rot_{i}"] = q[:, i]
  for i, k in enumerate(names):
    vertex[k] = cols[:, i]
  return vertex


def from_vertex_array(vertex, with_sh=True):
  """reference from_plydata (scene/io.py:70-117) -> dict of torch tensors (rotation back to normalised xyzw,
  alpha_logit (N,1) like the reference's Gaussians3D)."""
  get = lambda ks: torch.from_numpy(np.stack([np.asarray(vertex[k], dtype=np.float32) for k in ks], axis=-1))
  position = get(["x", "y", "z"])
  n = position.shape[0]
  names = sorted(vertex.dtype.names)
  if with_sh:
    n_sh = len([k for k in names if k.startswith("f_rest_") or k.startswith("f_dc_")]) // 3
    deg = int(np.sqrt(n_sh))
    assert deg * deg == n_sh, f"SH feature count must be square, got {n_sh}"
    dc = get([f"f_dc_{k}" for k in range(3)]).view(n, 3, 1)
    rest = get([f"f_rest_{k}" for k in range(3 * (n_sh - 1))]).view(n, 3, n_sh - 1) if n_sh > 1 else torch.zeros((n, 3, 0))
    feature = torch.cat([dc, rest], dim=2)
  else:
    feature = get([f"f_{k}" for k in range(len([k for k in names if k.startswith("f_")]))])
  rot = torch.nn.functional.normalize(get([f"rot_{k}" for k in range(4)]))
  out = dict(position=position, log_scaling=get([f"scale_{k}" for k in range(3)]), rotation=torch.roll(rot, -1, dims=(1,)),
             alpha_logit=get(["opacity"]))
  out["sh" if with_sh else "feature"] = feature
  return out


def write_ply(filename, vertex):
  header = ["ply", "format binary_little_endian 1.0", f"element vertex {vertex.shape[0]}"]
  inv = {"<f4": "float", "<f8": "double", "|u1": "uchar", "|i1": "char", "<i2": "short", "<u2": "ushort", "<i4": "int", "<u4": "uint"}
  for k in vertex.dtype.names:
    header.append(f"property {inv[vertex.dtype[k].newbyteorder('<').str if vertex.dtype[k].itemsize > 1 else vertex.dtype[k].str]} {k}")
  header.append("end_header")
  with open(filename, "wb") as f:
    f.write(("\n".join(header) + "\n").encode("ascii"))
    f.write(np.ascontiguousarray(vertex).tobytes())


def read_ply(filename):
  """the `vertex` element of a binary little-endian (or ascii) PLY with scalar properties."""
  with open(filename, "rb") as f:
    assert f.readline().strip() == b"ply", "not a PLY file"
    fmt, count, props, in_vertex = None, 0, [], False
    while True:
      line = f.readline()
      assert line, "unterminated PLY header"
      tok = line.decode("ascii").split()
      if not tok or tok[0] == "comment":
        continue
      if tok[0] == "format":
        fmt = tok[1]
      elif tok[0] == "element":
        in_vertex = tok[1] == "vertex"
        if in_vertex:
          count = int(tok[2])
      elif tok[0] == "property" and in_vertex:
        assert tok[1] != "list", "list properties are not part of the Gaussian PLY"
        props.append((tok[2], _PLY_TYPES[tok[1]]))
      elif tok[0] == "end_header":
        break
    dtype = np.dtype(props)
    if fmt == "binary_little_endian":
      return np.frombuffer(f.read(count * dtype.itemsize), dtype=dtype, count=count).copy()
    assert fmt == "ascii", f"unsupported PLY format {fmt}"
    rows = np.loadtxt(f, max_rows=count, ndmin=2)
    out = np.zeros(count, dtype=dtype)
    for i, (k, _) in enumerate(props):
      out[k] = rows[:, i]
    return out


def write_gaussians(filename, gaussians, with_sh=True):
  """reference write_gaussians (scene/io.py:119-124)."""
  write_ply(str(filename), to_vertex_array(gaussians, with_sh))


def read_gaussians(filename, with_sh=True):
  """reference read_gaussians (scene/io.py:128-132)."""
  return from_vertex_array(read_ply(str(filename)), with_sh)
