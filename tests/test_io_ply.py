"""The Gaussian PLY interchange (gsb200/io.py) against golden vectors produced by the reference's own
to_plydata / from_plydata (tests/golden/make_golden_io.py): field names, order and value conversions are pinned; the
file bytes follow the PLY specification (plyfile is absent) and are checked by round trip."""
import os

import numpy as np
import torch

from gsb200 import io as gio

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ply_io.npz"))


def _gaussians(kind):
  g = {k: G[k] for k in ("position", "log_scaling", "rotation", "alpha_logit")}
  g["sh" if kind == "sh" else "feature"] = G[kind]
  return g


def test_vertex_table_matches_reference():
  for kind in ("sh", "feature"):
    v = gio.to_vertex_array(_gaussians(kind), with_sh=(kind == "sh"))
    assert list(v.dtype.names) == [str(s) for s in G[f"{kind}_names"]]
    assert all(v.dtype[k] == np.dtype("<f4") for k in v.dtype.names)
    table = np.stack([v[k] for k in v.dtype.names], axis=1)
    np.testing.assert_allclose(table, G[f"{kind}_table"], rtol=1e-6, atol=1e-7)  # only the quaternion normalisation rounds


def test_file_round_trip_and_reference_read_back(tmp_path):
  for kind in ("sh", "feature"):
    path = tmp_path / f"{kind}.ply"
    gio.write_gaussians(path, _gaussians(kind), with_sh=(kind == "sh"))
    head = open(path, "rb").read(200).split(b"\n")
    assert head[0] == b"ply" and head[1] == b"format binary_little_endian 1.0" and head[2] == b"element vertex 57"
    assert head[3] == b"property float x"
    raw = gio.read_ply(str(path))
    assert np.array_equal(raw, gio.to_vertex_array(_gaussians(kind), with_sh=(kind == "sh")))
    back = gio.read_gaussians(path, with_sh=(kind == "sh"))
    for k in ("position", "rotation", "alpha_logit", "log_scaling"):
      np.testing.assert_allclose(back[k].numpy(), G[f"{kind}_back_{k}"], rtol=1e-6, atol=1e-7, err_msg=k)
    np.testing.assert_allclose(back["sh" if kind == "sh" else "feature"].numpy(), G[f"{kind}_back_feature"], rtol=0, atol=0)
    assert back["alpha_logit"].shape == (57, 1) and back["rotation"].shape == (57, 4)
    # xyzw in, normalised xyzw out
    q = torch.nn.functional.normalize(torch.from_numpy(G["rotation"]))
    np.testing.assert_allclose(back["rotation"].numpy(), q.numpy(), rtol=1e-6, atol=1e-7)


def test_ascii_ply_is_readable(tmp_path):
  path = tmp_path / "a.ply"
  v = gio.to_vertex_array(_gaussians("feature"), with_sh=False)[:3]
  with open(path, "w") as f:
    f.write("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 3\n")
    f.write("".join(f"property float {k}\n" for k in v.dtype.names) + "end_header\n")
    for row in v:
      f.write(" ".join(repr(float(x)) for x in row) + "\n")
  assert np.array_equal(gio.read_ply(str(path)), v)
