"""GPU parity of the colour MLP path (include/gsb200.h section 9) against the CPU oracle and against golden vectors
produced by the reference's own ColorModel (tests/golden/make_golden_color.py), including its fp16-autocast output —
the precision MLPScene.eval_colors really runs at (scene/mlp_scene.py:362)."""
import os

import numpy as np
import pytest
import torch

from gsb200 import ops
from gpu_util import cu, npy
from oracle import gs_oracle as go

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "color_mlp.npz"))
WEIGHTS = {k: G[k] for k in ops.COLOR_WEIGHTS}
# a colour is sigmoid * exp(intensity): relative error of a few fp16 ulps (2^-11 each) through two fp16 layers
TOL_REL, TOL_FLOOR = 4e-3, 0.25


def _close(ours, ref, rel=TOL_REL):
  err = np.abs(ours - ref)
  bound = rel * np.maximum(np.abs(ref), TOL_FLOOR)
  assert (err <= bound).all(), (err.max(), np.abs(ref)[err.argmax() // ref.shape[1]])


def _model():
  return ops.color_model({k: cu(v) for k, v in WEIGHTS.items()})


def test_forward_matches_reference_autocast_and_oracle():
  m = _model()
  spec = torch.zeros((G["feature"].shape[0], 3), device="cuda")
  rgb, spec = ops.color_mlp_fwd(m, cu(G["feature"]), cu(G["position"]), G["cam_pos"], cu(G["glo"]), specular=spec)
  rgb, spec = npy(rgb), npy(spec)
  # the reference itself under torch.autocast(float16)
  _close(spec, G["specular_amp"])
  _close(rgb, G["total_amp"])
  # the oracle with fp16 roundings at the same places
  d, s = go.Oracle("f64").color_mlp_fwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"], half=True)
  _close(spec, s)
  _close(rgb, d + s)
  # and the exact (float64) model: fp16 layers cost at most ~1 %
  _close(rgb, G["diffuse_f64"] + G["specular_f64"], rel=2e-2)


def test_specular_weight_and_in_view_rows():
  m = _model()
  n = G["feature"].shape[0]
  iv = torch.from_numpy((np.arange(n) % 3 != 0).astype(np.float32)).cuda()
  rgb = torch.full((n, 3), -7.0, device="cuda")
  spec = torch.full((n, 3), -7.0, device="cuda")
  ops.color_mlp_fwd(m, cu(G["feature"]), cu(G["position"]), G["cam_pos"], cu(G["glo"]), in_view=iv, specular_weight=0.25,
                    rgb=rgb, specular=spec)
  off = (iv == 0)
  assert (rgb[off] == -7.0).all() and (spec[off] == -7.0).all(), "rows out of view are left untouched"
  d, s = go.Oracle("f64").color_mlp_fwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"], half=True)
  on = npy(iv) > 0
  _close(npy(rgb)[on], (d + 0.25 * s)[on])


@pytest.mark.parametrize("n,p_view", [(1, 1.0), (31, 1.0), (70001, 0.55), (5000, 0.0)])
def test_forward_ragged_sizes_match_oracle(n, p_view):
  rng = np.random.default_rng(n)
  feature = rng.normal(0, 1.5, (n, 16)).astype(np.float32)
  position = rng.normal(0, 4, (n, 3)).astype(np.float32)
  cam = np.array([0.5, 0.25, -3.0], np.float32)
  glo = rng.normal(0, 0.7, 32).astype(np.float32)
  iv = (rng.uniform(0, 1, n) < p_view).astype(np.float32) * rng.uniform(0.1, 1.0, n).astype(np.float32)
  rgb, _ = ops.color_mlp_fwd(_model(), cu(feature), cu(position), cam, cu(glo), in_view=cu(iv))
  d, s = go.Oracle("f32").color_mlp_fwd(WEIGHTS, feature, position, cam, glo, in_view=iv, half=True)
  assert not npy(rgb)[iv == 0].any()
  _close(npy(rgb), d + s)


def test_unsupported_configuration_is_refused():
  from gsb200 import lib as L
  bad = ops.color_model({k: cu(v) for k, v in WEIGHTS.items()}, hidden=64)
  with pytest.raises(L.GsError):
    ops.color_mlp_fwd(bad, cu(G["feature"]), cu(G["position"]), G["cam_pos"], cu(G["glo"]))
