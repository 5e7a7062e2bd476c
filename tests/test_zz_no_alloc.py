"""gs_visible_indexes_static (include/gsb200.h): the allocation-free variant of gs_visible_indexes, whose scratch is a
file-scope device array inside libgsb200.so.  It has no counterpart in the reference;
the checker is the existing entry point gs_visible_indexes, itself parity-tested against the oracle's
`visible.nonzero()` in tests/test_gpu_optim_densify.py.  Integer work: bit-exact.

Named test_zz_* so that pytest collects it last.  Passes alone in a fresh process (the
first call looks the scratch address up) and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

from gsb200 import lib as L

pytestmark = pytest.mark.gpu


def _static(visible, indexes, count):
  """the new entry point, called through the C ABI: inputs, outputs and the stream, nothing else"""
  n = visible.shape[0]
  return L.load().gs_visible_indexes_static(n, L.ptr(visible), L.ptr(indexes), L.ptr(count), L.stream())


def _existing(visible):
  n = visible.shape[0]
  lib = L.load()
  indexes = torch.full((max(n, 1),), -1, device="cuda", dtype=torch.int32)
  count = torch.full((1,), -1, device="cuda", dtype=torch.int32)
  nbytes = lib.gs_visible_indexes_workspace_bytes(n)
  ws = torch.empty((nbytes,), device="cuda", dtype=torch.uint8)
  L.check(lib.gs_visible_indexes(n, L.ptr(visible), L.ptr(indexes), L.ptr(count), L.ptr(ws), nbytes, L.stream()))
  return indexes, count


def _visible(n, fraction, seed):
  rng = np.random.default_rng(seed)
  v = rng.uniform(0.01, 1.0, n).astype(np.float32)
  v[rng.uniform(size=n) >= fraction] = 0.0
  v[rng.uniform(size=n) < 0.01] = -1.0  # visible > 0 is the test, not != 0
  return torch.from_numpy(v).cuda()


def _check(n, fraction, seed):
  vis = _visible(n, fraction, seed)
  want_idx, want_count = _existing(vis)
  indexes = torch.full((max(n, 1),), -1, device="cuda", dtype=torch.int32)
  count = torch.full((1,), -1, device="cuda", dtype=torch.int32)
  L.check(_static(vis, indexes, count))
  torch.cuda.synchronize()
  k = int(want_count.item())
  assert int(count.item()) == k
  assert k == int((vis > 0).sum().item())
  assert torch.equal(indexes[:k], want_idx[:k])
  assert bool((indexes[k:] == -1).all())  # nothing written past the count
  return k


def test_two_calls_in_a_row_match_the_existing_entry_point():
  """the second call reuses the array the first one left dirty (ticket, status words): more tiles first, then fewer,
  then more again, with different contents each time"""
  k1 = _check(300_001, 0.30, seed=1)
  k2 = _check(70_003, 0.02, seed=2)
  k3 = _check(300_001, 0.90, seed=3)
  assert k1 > 0 and k2 > 0 and k3 > k1


@pytest.mark.parametrize("n", [1, 7, 2047, 2048, 2049, 4096 * 5 + 3])
def test_ragged_sizes_and_unaligned_input(n):
  _check(n, 0.5, seed=n)
  # a `visible` plane that does not start on a 16-byte boundary (a row slice of a larger plane)
  big = _visible(n + 3, 0.5, seed=n + 1)
  vis = big[3:]
  want_idx, want_count = _existing(vis.clone())
  indexes = torch.full((n,), -1, device="cuda", dtype=torch.int32)
  count = torch.zeros((1,), device="cuda", dtype=torch.int32)
  L.check(_static(vis, indexes, count))
  k = int(want_count.item())
  assert int(count.item()) == k and torch.equal(indexes[:k], want_idx[:k])


def test_empty_and_none_visible():
  count = torch.full((1,), 5, device="cuda", dtype=torch.int32)
  L.check(L.load().gs_visible_indexes_static(0, None, None, L.ptr(count), L.stream()))
  assert int(count.item()) == 0
  vis = torch.zeros((10_000,), device="cuda")
  indexes = torch.full((10_000,), -1, device="cuda", dtype=torch.int32)
  L.check(_static(vis, indexes, count))
  assert int(count.item()) == 0 and bool((indexes == -1).all())


def test_full_size_at_the_limit_and_rejection_beyond_it():
  lib = L.load()
  limit = int(lib.gs_visible_indexes_static_max_rows())
  assert limit == 1 << 24
  _check(limit, 0.05, seed=9)
  # one row more: refused with a clear error before anything is touched (the pointers are never dereferenced)
  count = torch.full((1,), 7, device="cuda", dtype=torch.int32)
  dummy = torch.zeros((16,), device="cuda")
  rc = lib.gs_visible_indexes_static(limit + 1, L.ptr(dummy), L.ptr(dummy), L.ptr(count), L.stream())
  assert rc == -1  # GS_ERR_INVALID
  msg = lib.gs_last_error().decode()
  assert "GS_VISIBLE_INDEXES_STATIC_MAX_ROWS" in msg and str(limit) in msg
  with pytest.raises(L.GsError):
    L.check(lib.gs_visible_indexes_static(-1, L.ptr(dummy), L.ptr(dummy), L.ptr(count), L.stream()))
  torch.cuda.synchronize()
  assert int(count.item()) == 7
  _check(4097, 0.5, seed=10)  # and the entry point still works afterwards


def test_no_device_allocation_during_the_call():
  """the driver's free-memory figure does not move across 20 calls (torch's caching allocator is not involved: inputs and
  outputs are allocated before the first reading)"""
  n = 1_000_000
  vis = _visible(n, 0.3, seed=4)
  indexes = torch.empty((n,), device="cuda", dtype=torch.int32)
  count = torch.zeros((1,), device="cuda", dtype=torch.int32)
  L.check(_static(vis, indexes, count))  # set-up call: the scratch address is looked up here
  torch.cuda.synchronize()
  free0, _ = torch.cuda.mem_get_info()
  reserved0 = torch.cuda.memory_reserved()
  for _ in range(20):
    L.check(_static(vis, indexes, count))
  torch.cuda.synchronize()
  free1, _ = torch.cuda.mem_get_info()
  assert free1 >= free0 and torch.cuda.memory_reserved() == reserved0  # nothing was taken from the device
  assert int(count.item()) == int((vis > 0).sum().item())
