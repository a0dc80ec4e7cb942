"""Multi-rank host logic on CPU: world size 2, gloo backend.

Covers the block partition of columns, the spectral split (chunks share end points, carry
a cut-off halo, and their sub-grids have exactly the expected number of points when built
by the host library), and the one collective of the path -- the SUM all-reduce of partial
broadband integrals -- against the unsplit trapezoid sum (framework/src/driver.c:302-319).
No CUDA call is made here; the GPU counterpart is tests/test_gpu_parity.py.
"""
import os
import socket

import numpy as np
import pytest

from grtcode_b200 import parallel as P


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_column_shards_cover_every_column_once():
    for cols, world in ((1800, 8), (1800, 7), (65341, 8), (5, 8), (1, 1)):
        seen = []
        for r in range(world):
            seen.extend(P.shard_columns(cols, world, r))
        assert seen == list(range(cols))
        sizes = [len(P.shard_columns(cols, world, r)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        P.shard_columns(10, 2, 2)


@pytest.mark.parametrize("w0,wn,dw,world", [(1., 50000., 0.01, 8), (1., 3250., 0.1, 2),
                                            (1., 3000., 0.1, 3), (500., 700., 0.1, 4)])
def test_spectral_chunks_tile_the_grid(w0, wn, dw, world):
    n = P.grid_size(w0, wn, dw)
    chunks = P.split_spectral_range(w0, wn, dw, world)
    assert chunks[0].first == 0 and chunks[-1].last == n - 1
    halo = int(np.ceil(np.float32(25.)/dw)) + 1
    for a, b in zip(chunks[:-1], chunks[1:]):
        assert a.last == b.first                      # shared seam point
    for c in chunks:
        assert c.lo == max(c.first - halo, 0) and c.hi == min(c.last + halo, n - 1)
        first, last = c.window
        assert 0 <= first < last <= c.n - 1
        assert abs(c.w_lo - (w0 + c.lo*dw)) < 1e-9
    sizes = [c.last - c.first for c in chunks]
    assert sum(sizes) == n - 1 and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        P.split_spectral_range(1., 1.5, 0.1, 8)


def test_cost_balanced_chunks(tmp_path):
    """With a per-interval cost the cuts equalise the summed cost; the chunks still tile the
    grid, share their seam points and carry the same halo.  A uniform cost reproduces the
    equal-interval split (to the rounding of a cut)."""
    from grtcode_b200 import synth
    w0, wn, dw, world = 1., 3250., 0.1, 4
    n = P.grid_size(w0, wn, dw)
    same = P.split_spectral_range(w0, wn, dw, world, np.ones(n - 1))
    plain = P.split_spectral_range(w0, wn, dw, world)
    assert all(abs(a.first - b.first) <= 1 and abs(a.last - b.last) <= 1
               for a, b in zip(same, plain))
    # a line list crowded into 600-800 cm-1, written and read back as HITRAN records
    rng = np.random.default_rng(3)
    sets = [synth.make_lines(synth.CO2, 4000, 600., 800., rng),
            synth.make_lines(synth.H2O, 500, w0, wn, rng)]
    path = str(tmp_path/"lines.par")
    synth.write_hitran(path, sets)
    v = P.line_centres(path)
    assert v.size == 4500 and v.min() >= w0 and v.max() <= wn
    only = P.line_centres(path, molecules=(synth.H2O,))
    assert only.size == 500
    cost = P.interval_cost(w0, wn, dw, v, doppler_slope=0.)
    assert cost.shape == (n - 1,) and np.all(cost > 0.)
    # total line cost is conserved by the box filter away from the grid ends
    assert abs((cost - 2.*dw).sum() - v.size) < 0.02*v.size
    chunks = P.split_spectral_range(w0, wn, dw, world, cost)
    assert chunks[0].first == 0 and chunks[-1].last == n - 1
    for a, b in zip(chunks[:-1], chunks[1:]):
        assert a.last == b.first and a.first < a.last
    shares = np.array([cost[c.first:c.last].sum() for c in chunks])/cost.sum()
    assert np.all(np.abs(shares - 1./world) < 0.02), shares
    plain_shares = np.array([cost[c.first:c.last].sum() for c in plain])/cost.sum()
    assert plain_shares.max() > 0.5                   # what equal point counts would give
    with pytest.raises(ValueError):
        P.split_spectral_range(w0, wn, dw, world, np.ones(n))
    # degenerate: all the cost in one interval still leaves every rank an interval
    spike = np.zeros(n - 1)
    spike[10] = 1.
    for c in P.split_spectral_range(w0, wn, dw, world, spike):
        assert c.last > c.first


def test_chunk_grids_have_the_expected_size_in_the_host_library(built):
    """create_spectral_grid (spectral_grid.c:61) on (w_lo, wn_arg, dw) gives hi - lo + 1
    points for every chunk of the config-4 split, and w(i) lands on the global grid."""
    api = built
    api.load()
    for w0, wn, dw, world in ((1., 50000., 0.01, 8), (1., 3250., 0.1, 4)):
        for c in P.split_spectral_range(w0, wn, dw, world):
            g = api.spectral_grid(c.w_lo, c.wn_arg, dw)
            assert int(g.n) == c.n
            assert abs((g.w0 + (c.n - 1)*dw) - (w0 + c.hi*dw)) < 1e-6*dw


def _planck_like(w):
    return w**3/np.expm1(w/600.) + 50.*np.exp(-0.5*((w - 667.)/8.)**2)


def _worker(rank: int, world: int, port: int, tmp: str):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        w0, wn, dw, levels = 1., 3250., 0.1, 7
        n = P.grid_size(w0, wn, dw)
        ch = P.split_spectral_range(w0, wn, dw, world)[rank]
        # synthetic spectral fluxes on the rank's own (halo-carrying) grid
        w = ch.w_lo + np.arange(ch.n)*dw
        scale = 1. + 0.1*np.arange(levels)[:, None]
        up_spec = scale*_planck_like(w)[None, :]
        dn_spec = 0.5*scale*_planck_like(w)[None, :]
        first, last = ch.window
        up, dn = P.reduce_broadband(P.trapezoid(up_spec[:, first:last + 1], dw),
                                    P.trapezoid(dn_spec[:, first:last + 1], dw))
        wg = w0 + np.arange(n)*dw
        want_up = P.trapezoid(scale*_planck_like(wg)[None, :], dw)
        np.testing.assert_allclose(up, want_up, rtol=1e-12)
        np.testing.assert_allclose(dn, 0.5*want_up, rtol=1e-12)

        # column sharding: each rank fills its block, rank 0 gathers all 11 columns
        cols = 11
        mine = P.shard_columns(cols, world, rank)
        local = np.array([[c, 10.*c + 1.] for c in mine], dtype=np.float64).reshape(len(mine), 2)
        allc = P.gather_columns(local, cols)
        if rank == 0:
            assert allc.shape == (cols, 2)
            np.testing.assert_array_equal(allc[:, 0], np.arange(cols))
            np.testing.assert_array_equal(allc[:, 1], 10.*np.arange(cols) + 1.)
        else:
            assert allc is None
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_ranks_reduce_and_gather_over_gloo(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_single_process_reduce_is_the_identity():
    up, dn = P.reduce_broadband(np.arange(4.), np.arange(4.) + 1.)
    np.testing.assert_array_equal(up, np.arange(4.))
    np.testing.assert_array_equal(dn, np.arange(4.) + 1.)


def test_measured_costs_move_the_cuts_towards_the_slow_ranks():
    """parallel.measured_costs: chunk sums equal the measured seconds, and cutting by the
    result gives the slow rank fewer intervals."""
    from grtcode_b200 import parallel as P
    w0, wn, dw, world = 1., 1001., 0.5, 4
    chunks = P.split_spectral_range(w0, wn, dw, world)
    n_int = P.grid_size(w0, wn, dw) - 1
    secs = [1., 1., 3., 1.]
    cost = P.measured_costs(None, chunks, secs)
    assert cost.shape == (n_int,)
    for ch, t in zip(chunks, secs):
        assert abs(cost[ch.first:ch.last].sum() - t) < 1e-12
    again = P.split_spectral_range(w0, wn, dw, world, cost)
    width = lambda c: c.last - c.first
    assert sum(width(c) for c in again) == n_int
    # every rank now carries the same measured cost: 6/4 s
    for c in again:
        assert abs(cost[c.first:c.last].sum() - 1.5) < 2.*cost.max()
    # the slow interval range [chunks[2].first, chunks[2].last) is shared by more ranks
    assert sum(1 for c in again if c.first < chunks[2].last and c.last > chunks[2].first) >= 2
    shaped = P.measured_costs(np.linspace(1., 2., n_int), chunks, secs)
    seg = shaped[chunks[0].first:chunks[0].last]
    assert seg[-1] > seg[0]            # the model's shape is kept inside a chunk


def test_halo_covers_the_pressure_shift(tmp_path):
    """A line whose unshifted centre lies just outside a rank's cut-off halo but whose shifted
    centre reaches an owned point must be on the rank's grid: the halo grows by
    ceil(max|delta| p/dw) points."""
    from grtcode_b200 import parallel as P, synth
    plain = P.split_spectral_range(1., 1001., 0.01, 4)
    wide = P.split_spectral_range(1., 1001., 0.01, 4, shift=0.031)
    for a, b in zip(plain, wide):
        assert (a.first, a.last) == (b.first, b.last)
        if a.lo > 0:
            assert a.lo - b.lo == 4
        if a.hi < 100000:
            assert b.hi - a.hi == 4
    ds = synth.make_dataset(str(tmp_path), 100., 200., {1: 50, 2: 50}, seed=1)
    d = P.max_pressure_shift(ds.hitran)
    assert 0.01 < d <= 0.02        # synth draws delta from U(-0.02, 0.015)
    assert P.max_pressure_shift(ds.hitran, molecules=(7,)) == 0.
