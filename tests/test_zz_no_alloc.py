"""mpmk_particle_volume_noalloc (include/mpmk.h; no counterpart in the reference -- added
for callers that cannot allocate): the particle-volume kernels of internel_filling/filling.py:12-24 with their
per-cell scratch in a file-scope device array inside the library.  Called through the C ABI and compared with the existing
entry point mpmk_particle_volume (caller-provided, caller-zeroed scratch) on two calls in a row -- the second on other
positions, so that counts left behind by the first would show.  The two launch the same kernels on the same integer counts:
the volumes must agree bit for bit (tolerance 0, inside the 1e-6 the G2P-level tests use).
Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_GRID, EXTENT = 50, 2.0


def _positions(n, seed, lo, hi):
    rng = np.random.default_rng(seed)
    # clustered, so that cells hold from one to hundreds of particles
    centres = rng.uniform(lo + 0.1, hi - 0.1, size=(12, 3))
    p = centres[rng.integers(0, 12, n)] + rng.standard_normal((n, 3)) * 0.05
    return np.clip(p, lo, hi).astype(np.float32)


def _both(ctx, pos, dev):
    import torch
    n = pos.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    dx = EXTENT / N_GRID
    cells = torch.zeros(N_GRID ** 3, dtype=torch.int32, device=dev)
    want = torch.empty(n, dtype=torch.float32, device=dev)
    ctx.call("mpmk_particle_volume", n, pos.data_ptr(), float(dx), cells.data_ptr(), want.data_ptr(), st)
    got = torch.full((n,), -1.0, dtype=torch.float32, device=dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    alloc0 = torch.cuda.memory_allocated(dev)
    ctx.call("mpmk_particle_volume_noalloc", n, pos.data_ptr(), float(dx), got.data_ptr(), st)
    torch.cuda.synchronize()
    d_free, d_alloc = free0 - torch.cuda.mem_get_info(dev)[0], torch.cuda.memory_allocated(dev) - alloc0   # before anything else runs
    return got.cpu().numpy(), want.cpu().numpy(), cells.cpu().numpy(), d_free, d_alloc


def test_particle_volume_noalloc_matches_existing_entry_point(cuda_dev):
    import torch
    from mpm_solver import _native
    ctx = _native.Context(cuda_dev.index or 0, N_GRID, EXTENT, 0)
    try:
        dx = EXTENT / N_GRID
        deltas = []
        for call, (n, seed, lo, hi) in enumerate(((60000, 1, 0.5, 1.5), (35000, 2, 0.2, 1.1))):
            pos = torch.from_numpy(_positions(n, seed, lo, hi)).to(cuda_dev)
            got, want, cells, d_free, d_alloc = _both(ctx, pos, cuda_dev)
            assert np.array_equal(got, want), (call, float(np.abs(got - want).max()))
            # and against the definition: vol = dx^3 / (particles in the same cell)
            c = np.floor(pos.cpu().numpy().astype(np.float32) / np.float32(dx)).astype(np.int64)
            flat = (c[:, 0] * N_GRID + c[:, 1]) * N_GRID + c[:, 2]
            cnt = np.bincount(flat, minlength=N_GRID ** 3)
            assert np.array_equal(cnt, cells) and cnt.max() > 50 and (cnt == 1).any()
            assert np.allclose(got, np.float32(dx) ** 3 / cnt[flat], rtol=1e-6, atol=0.0)
            assert d_alloc == 0                     # nothing went through torch's allocator
            deltas.append(d_free)
        assert deltas[1] == 0, deltas              # and the driver's free memory did not move on the second call
    finally:
        ctx.close()


def test_particle_volume_noalloc_rejects_grids_beyond_the_builtin_scratch(cuda_dev):
    import torch
    from mpm_solver import _native
    ctx = _native.Context(cuda_dev.index or 0, 256, EXTENT, 0)
    try:
        pos = torch.rand(100, 3, device=cuda_dev) * EXTENT
        vol = torch.empty(100, device=cuda_dev)
        with pytest.raises(_native.MpmkError, match="exceeds the built-in scratch"):
            ctx.call("mpmk_particle_volume_noalloc", 100, pos.data_ptr(), EXTENT / 256, vol.data_ptr(), torch.cuda.current_stream().cuda_stream)
    finally:
        ctx.close()
