"""lbm_population_sums_noalloc through the C ABI against lbm_diagnostics.

The entry point serves callers that may not allocate device memory (their own pool owns it all, or they run
where the device-wide wait inside cudaMalloc/cudaFree would stall somebody else's work): its scratch is a
fixed array inside the library, shared by all calls and never cleared.  So the test makes TWO calls in a row
on different data -- the second must not be affected by what the first left behind -- and compares each with
the context-based lbm_diagnostics on the same populations.  The buffers handed in are allocated here by
torch; the entry point itself allocates nothing.  Collected last on purpose (file name), and it must pass
both alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12   # relative; same bound as the diagnostics test in test_gpu_parity.py (summation order differs)


def _reference_sums(api, nx, ny, f_aos):
    ctx = api.Context(api.mrt_config("gs", nx, ny))     # periodic model: every node counts
    ctx.set_populations(0, f_aos)
    d = ctx.diagnostics()
    ctx.close()
    return d


def test_noalloc_sums_match_diagnostics_twice():
    import torch
    from latticeboltzmannmethod_b200 import api
    nx, ny = 300, 217                                   # 65 100 nodes: several 8192-node blocks plus a ragged tail
    n = nx * ny
    rng = np.random.default_rng(2024)
    out = torch.zeros(6, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        f = rng.uniform(0.01, 0.2, (n, 9)) * (1.0 + call)          # second call: different data, different sums
        if call == 1:
            f[12345, 4] = np.nan                                     # and a NaN that the first call did not have
        want = _reference_sums(api, nx, ny, f)
        soa = torch.from_numpy(np.ascontiguousarray(f.T)).cuda()     # [k][node], stride_k = n
        api.population_sums_noalloc(soa.data_ptr(), 0, 9, n, n, stream, out.data_ptr())
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert bool(got[5] != 0.0) == want["has_nan"] == (call == 1)
        if call == 0:
            scale = abs(want["mass"])
            assert abs(got[0] - want["mass"]) <= TOL * scale
            assert abs(got[1] - want["momentum"][0]) <= TOL * scale
            assert abs(got[2] - want["momentum"][1]) <= TOL * scale
            assert got[3] == 0.0 and got[4] == 0.0
        else:
            assert np.isnan(got[0]) and np.isnan(want["mass"])       # the NaN propagates into the mass in both
            assert got[5] == 1.0


def test_noalloc_d3q19_with_g_and_a_nan_in_g_twice():
    """The 19-direction table and the g accumulation: two calls in a row on different data, the second with a NaN
    placed in g only (f stays clean, so the f sums stay finite and comparable).  Checked against float64 numpy sums
    of the same arrays (summation order differs: relative tolerance)."""
    import torch
    from latticeboltzmannmethod_b200 import api
    cx = np.array([0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1], dtype=np.float64)
    cy = np.array([0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0], dtype=np.float64)
    cz = np.array([0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1], dtype=np.float64)
    n, stride = 50_021, 50_048                           # ragged tail; directions padded apart (stride_k > n_nodes)
    rng = np.random.default_rng(77)
    out = torch.zeros(6, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        f = rng.uniform(0.01, 0.2, (19, stride)) * (1.0 + 2 * call)
        g = rng.uniform(-0.1, 0.1, (19, stride)) * (1.0 + call)
        if call == 1:
            g[7, 31_000] = np.nan
        fd, gd = torch.from_numpy(f).cuda(), torch.from_numpy(g).cuda()
        api.population_sums_noalloc(fd.data_ptr(), gd.data_ptr(), 19, n, stride, stream, out.data_ptr())
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        fn = f[:, :n]
        scale = abs(fn.sum())
        assert abs(got[0] - fn.sum()) <= TOL * scale
        for c, tab in zip((1, 2, 3), (cx, cy, cz)):
            assert abs(got[c] - (tab[:, None] * fn).sum()) <= TOL * scale, c
        if call == 0:
            assert abs(got[4] - g[:, :n].sum()) <= TOL * np.abs(g[:, :n]).sum()
            assert got[5] == 0.0
        else:
            assert np.isnan(got[4]) and got[5] == 1.0    # one NaN, counted once, and it poisons the g sum only


def test_noalloc_rejects_sizes_beyond_the_scratch():
    import torch
    from latticeboltzmannmethod_b200 import api
    out = torch.zeros(6, dtype=torch.float64, device="cuda")
    small = torch.zeros(9 * 16, dtype=torch.float64, device="cuda")
    too_many = 16384 * 8192 + 1
    with pytest.raises(api.LbmError) as e:
        api.population_sums_noalloc(small.data_ptr(), 0, 9, too_many, too_many, 0, out.data_ptr())
    assert "exceed" in str(e.value)
    with pytest.raises(api.LbmError):
        api.population_sums_noalloc(small.data_ptr(), 0, 7, 16, 16, 0, out.data_ptr())
