"""Pins the colour-MLP oracle (oracle/gs_oracle.c section 9) to golden vectors produced by the reference's own
ColorModel / rsh code (tests/golden/make_golden_color.py): real spherical harmonics, fp32 forward, the fp16-autocast
forward MLPScene.eval_colors really runs, and float64 autograd gradients."""
import os

import numpy as np
import pytest

from oracle import gs_oracle as go

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "color_mlp.npz"))
WEIGHTS = {k: G[k] for k in go.Oracle.COLOR_WEIGHTS}


@pytest.fixture(scope="module")
def orc():
  return go.Oracle("f64")


def test_rsh_matches_reference_closed_forms(orc):
  np.testing.assert_allclose(orc.rsh(5, G["dirs"]), G["rsh5"], rtol=1e-6, atol=1e-6)  # dirs were stored as float32
  np.testing.assert_allclose(orc.rsh(3, G["dirs"]), G["rsh3"], rtol=1e-6, atol=1e-6)


def test_forward_matches_reference_fp64_and_fp32(orc):
  d, s = orc.color_mlp_fwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"])
  np.testing.assert_allclose(d, G["diffuse_f64"], rtol=1e-9, atol=1e-12)
  np.testing.assert_allclose(s, G["specular_f64"], rtol=1e-9, atol=1e-12)
  np.testing.assert_allclose(d, G["diffuse_f32"], rtol=2e-5, atol=1e-6)
  np.testing.assert_allclose(s, G["specular_f32"], rtol=2e-5, atol=1e-6)
  assert d.max() > 1.5 and s.max() > 0.2 and d.min() < 0.05, "the golden model leaves the linear range"


def test_half_rounded_forward_tracks_the_autocast_reference(orc):
  """fp16 roundings where autocast puts them on the Linear layers; the reference additionally rounds the activation
  arithmetic (exp, sigmoid, products) to fp16, which the oracle evaluates exactly: a few fp16 ulps of the output."""
  assert str(G["amp_dtype"]) == "torch.float16"
  d, s = orc.color_mlp_fwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"], half=True)
  for ours, ref in ((d, G["diffuse_amp"]), (s, G["specular_amp"]), (d + s, G["total_amp"])):
    err = np.abs(ours - ref)
    assert (err <= 4e-3 * np.maximum(np.abs(ref), 0.25)).all(), err.max()
  # and the rounding is not a no-op: the fp16 path differs measurably from fp64
  assert np.abs(d - G["diffuse_f64"]).max() > 1e-4


def test_in_view_rows_only(orc):
  iv = (np.arange(G["feature"].shape[0]) % 3 != 0).astype(np.float64)
  d, s = orc.color_mlp_fwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"], in_view=iv)
  assert not d[iv == 0].any() and not s[iv == 0].any()
  np.testing.assert_allclose(d[iv > 0], G["diffuse_f64"][iv > 0], rtol=1e-9)


def test_backward_matches_reference_autograd(orc):
  df, dg, dW = orc.color_mlp_bwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"], G["cd"], G["cs"])
  np.testing.assert_allclose(df, G["d_feature"], rtol=1e-7, atol=1e-10)
  np.testing.assert_allclose(dg, G["d_glo"], rtol=1e-7, atol=1e-9)
  for k in go.Oracle.COLOR_WEIGHTS:
    np.testing.assert_allclose(dW[k], G["d_" + k], rtol=1e-7, atol=1e-9, err_msg=k)
  # row subset: gradients of the selected rows only
  w = (np.arange(G["feature"].shape[0]) % 2).astype(np.float64)
  df2, _, _ = orc.color_mlp_bwd(WEIGHTS, G["feature"], G["position"], G["cam_pos"], G["glo"], G["cd"], G["cs"], row_w=w)
  assert not df2[w == 0].any()
  np.testing.assert_allclose(df2[w > 0], G["d_feature"][w > 0], rtol=1e-7, atol=1e-10)
