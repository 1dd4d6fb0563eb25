"""titan_b200.lattice (the product-side direct-to-array builder) against the oracle's restatement of
Lattice::Lattice (itself pinned to the reference's builder in test_oracle_pin.py): bit-exact topology."""
import numpy as np
import pytest

import oracle
from titan_b200 import lattice


@pytest.mark.parametrize("shape", [(5, 5, 5), (1, 1, 1), (1, 5, 2), (3, 4, 6), (10, 10, 10), (2, 1, 7), (13, 9, 11)])
def test_bit_exact_against_oracle(shape):
    c, d = (0.1, -0.2, 3.3), (2.0, 3.0, 4.0)
    o = oracle.lattice(c, d, *shape)
    g = lattice.build_lattice(c, d, *shape)
    assert lattice.lattice_counts(*shape) == (o["pos"].shape[0], o["left"].shape[0])
    assert (g["pos"].view(np.uint64) == o["pos"].view(np.uint64)).all()
    assert (g["left"] == o["left"]).all() and (g["right"] == o["right"]).all()
    assert (g["rest"].view(np.uint64) == o["rest"].view(np.uint64)).all()


def test_chunked_emission_order_is_chunk_independent():
    a = lattice.lattice_springs(9, 4, 5, i_planes_per_chunk=2)
    b = lattice.lattice_springs(9, 4, 5, i_planes_per_chunk=100)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


def test_survey_sizes():
    assert lattice.lattice_counts(100, 100, 100) == (1000000, 12731796)
    assert lattice.lattice_counts(300, 300, 300) == (27000000, 348575396)
    assert lattice.lattice_counts(200, 200, 200) == (8000000, 102923596)


@pytest.mark.parametrize("shape", [(38, 60, 60), (37, 60, 60), (12, 24, 24), (11, 24, 24), (100, 100, 100), (7, 3, 50),
                                   (94, 40, 40)])
def test_brick_hint_tiles_are_valid_and_layered_shapes_need_fewer_tiles(shape, monkeypatch):
    """tb_set_tile_hint input for lattices (titan_b200.lattice.brick_hint): ids dense from 0, at most 128 masses per id,
    every id a box of lattice cells; cutting the i axis into layers of differently shaped bricks never needs more
    tiles than the single best shape (8-GPU slabs of 37/38 planes: DESIGN.md section 5)."""
    nx, ny, nz = shape
    h = lattice.brick_hint(nx, ny, nz).reshape(nx, ny, nz)
    monkeypatch.setenv("TB_BRICK_LAYERS", "0")
    h0 = lattice.brick_hint(nx, ny, nz).reshape(nx, ny, nz)
    for ids in (h, h0):
        counts = np.bincount(ids.ravel())
        assert counts.min() >= 1 and counts.max() <= 128 and counts.size == ids.max() + 1
    assert h.max() <= h0.max()
    # every tile is an axis-aligned box: its bounding box holds exactly its masses
    ii, jj, kk = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    nt = int(h.max()) + 1
    vol = np.ones(nt, dtype=np.int64)
    for c in (ii, jj, kk):
        lo = np.full(nt, 1 << 30); hi = np.full(nt, -1)
        np.minimum.at(lo, h.ravel(), c.ravel()); np.maximum.at(hi, h.ravel(), c.ravel())
        vol *= hi - lo + 1
    assert (vol == np.bincount(h.ravel())).all()
    if nx % 5 == 0 and ny % 5 == 0 and nz % 5 == 0 and min(shape) >= 5:
        assert (h == h0).all()                 # nothing to gain: the numbering of the single shape is kept
