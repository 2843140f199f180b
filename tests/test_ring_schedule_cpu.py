"""Host check of the ring ordering of the Jacobi eigensolvers (csrc/eigh_ring.cuh, RingCols::phases; restated in
tests/dev/proto/ring_schedule.py): for every processor count the kernels use, every pair of columns meets in a sweep,
and the sweep has the number of rounds the kernels' documentation states."""
import importlib.util
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location('ring_schedule', os.path.join(HERE, 'dev', 'proto', 'ring_schedule.py'))
ring_schedule = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ring_schedule)


@pytest.mark.parametrize('K0,rounds', [(4, 7), (8, 15), (16, 31), (10, 20), (12, 24), (14, 28)])
def test_every_pair_meets(K0, rounds):
    n, missing, most = ring_schedule.sweep(K0)
    assert not missing
    assert n == rounds
    assert most == (1 if K0 & (K0 - 1) == 0 else 2)      # powers of two: exactly once; else the byes of the round robin twice
