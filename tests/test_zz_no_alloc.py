"""GPU test of txf_moments_update_noalloc, kept in a file pytest collects last.

That entry point is for hosts that embed the library as a plug-in and are handed input/output
buffers but no device allocator (a Beam/TFX worker, a serving process with a fixed memory
budget): its scratch is an array inside the library, so it must give exactly what the
workspace-taking entry point gives -- here on two calls in a row, which share that array.
"""
import ctypes

import numpy as np
import pytest
import torch

from helpers import O

pytestmark = pytest.mark.gpu


def _moments(nat, name, cols, n, with_workspace):
    """Accumulate `cols` (list of (values, valid-or-None)) into a fresh state through entry point `name`."""
    ncols = len(cols)
    state = torch.zeros(ncols * 3, dtype=torch.float64, device="cuda")
    vals = nat.ptr_array([v.data_ptr() for v, _ in cols])
    valid = nat.ptr_array([None if m is None else m.data_ptr() for _, m in cols])
    dtype = nat.i32_array([nat.TXF_F32 if v.dtype == torch.float32 else nat.TXF_I64 for v, _ in cols])
    if with_workspace:
        ws = torch.empty(nat.query("txf_moments_workspace_bytes", ncols), dtype=torch.uint8, device="cuda")
        nat.call(name, vals, valid, dtype, ncols, n, state.data_ptr(), ws.data_ptr(), None)
    else:
        nat.call(name, vals, valid, dtype, ncols, n, state.data_ptr(), None)
    torch.cuda.synchronize()
    return state.cpu().numpy().reshape(ncols, 3)


def test_no_alloc_moments_match_workspace_entry_point_twice():
    import kubeflow_tfx_workshop_b200 as ktw
    nat = ktw._native
    rng = np.random.default_rng(42)
    for call, n in enumerate((1_000_003, 250_001)):  # two calls in a row share the built-in scratch
        f = torch.from_numpy(rng.normal(10 + call, 3, n).astype(np.float32)).cuda()
        g = torch.from_numpy((rng.random(n) * 700).astype(np.float32)).cuda()
        i = torch.from_numpy(rng.integers(0, 86_400, n)).cuda()
        bits = torch.from_numpy(np.packbits(rng.random(n) > 0.1, bitorder="little")).cuda()
        cols = [(f, None), (g, bits), (i, bits)]
        got = _moments(nat, "txf_moments_update_noalloc", cols, n, with_workspace=False)
        want = _moments(nat, "txf_moments_update", cols, n, with_workspace=True)
        # same kernels, same grid, same merge order: the two entry points agree to the last bit
        assert np.array_equal(got, want), (call, got, want)
        assert got[0, 0] == n
        # and, independently of the other entry point, what the oracle computes for the filled columns
        present = np.unpackbits(bits.cpu().numpy(), bitorder="little")[:n].astype(bool)
        for c, (v, m) in enumerate(cols):
            x = v.cpu().numpy().astype(np.float32)
            cnt, mean, m2 = O.moments(np.where(present, x, np.float32(0)) if m is not None else x)
            assert got[c, 0] == cnt
            assert abs(got[c, 1] - mean) <= 1e-12 * max(abs(mean), 1.0), (call, c)
            assert abs(got[c, 2] - m2) <= 1e-10 * max(abs(m2), 1.0), (call, c)
    # beyond the built-in scratch: a clear error instead of an allocation
    cols5 = [(torch.zeros(64, device="cuda"), None)] * 5
    with pytest.raises(nat.NativeError, match="at most 4 columns"):
        _moments(nat, "txf_moments_update_noalloc", cols5, 64, with_workspace=False)
