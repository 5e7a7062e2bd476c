"""CPU: the C-ABI library loads and exports every symbol include/gsb200.h declares; the binding covers them all;
the product path never touches oracle/ and has no CPU fallback.  (No compute calls: there is no GPU here.)"""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gsb200.h")
PKG = os.path.join(ROOT, "splat-trainer_b200")
LIB = os.path.join(PKG, "libgsb200.so")


def declared_functions():
  src = open(HEADER).read()
  src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
  return sorted(set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def built():
  if not os.path.exists(LIB):
    subprocess.run(["make", "-C", os.path.join(PKG, "csrc"), "-j8"], check=True, capture_output=True)
  return LIB


def test_header_declares_the_hot_path():
  names = declared_functions()
  for must in ("gs_project_fwd", "gs_project_bwd", "gs_tile_map", "gs_radix_sort_pairs", "gs_raster_fwd", "gs_raster_bwd",
               "gs_optim_step", "gs_accumulate_stats", "gs_select_topk_mask", "gs_compact_rows", "gs_split_positions"):
    assert must in names


def test_library_exports_every_declared_symbol(built):
  lib = ctypes.CDLL(built)
  for name in declared_functions():
    assert hasattr(lib, name), f"{name} declared in include/gsb200.h but not exported by libgsb200.so"


def test_binding_covers_every_declared_symbol(built):
  from gsb200 import lib as L
  assert sorted(L.SIGNATURES) == declared_functions()
  L.load()
  assert L.load().gs_version() == 100
  cfg = L.default_raster_config()
  assert cfg.tile_size == 16 and abs(cfg.alpha_threshold - 1 / 255) < 1e-9 and abs(cfg.blur_cov - 0.3) < 1e-7


def test_no_torch_in_the_abi(built):
  out = subprocess.run(["nm", "-D", "--defined-only", built], capture_output=True, text=True).stdout
  exported = [l.split()[-1] for l in out.splitlines() if " T " in l]
  assert exported and all(not re.search(r"torch|at::|c10", s) for s in exported)
  deps = subprocess.run(["ldd", built], capture_output=True, text=True).stdout
  libs = [l.split()[0] for l in deps.splitlines() if l.strip()]  # names only: load addresses are random hex
  assert not any(re.search(r"torch|c10", l) for l in libs)
  code = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)  # declarations only
  assert not re.search(r"torch|at::|Tensor", code)


def test_product_never_touches_the_oracle_and_has_no_cpu_fallback():
  for dirpath, _, files in os.walk(PKG):
    if "build" in dirpath:
      continue
    for f in files:
      if f.endswith((".py", ".cu", ".cuh", ".h")):
        text = open(os.path.join(dirpath, f)).read()
        assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f"{f} imports the oracle"
        assert "gs_oracle" not in text, f"{f} references the oracle"
  import torch
  from gsb200 import lib as L
  from gsb200 import ops
  with pytest.raises(L.GsError):
    ops.select_topk_mask(torch.zeros(8), 2, 0)  # CPU tensors are refused, not silently computed
  with pytest.raises(RuntimeError):
    L.load(os.path.join(PKG, "does_not_exist.so"))
