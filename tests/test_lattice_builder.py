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
