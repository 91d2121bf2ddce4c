"""gc_radius_of_gyration_noalloc: the allocation-free variant of gc_radius_of_gyration (static
device scratch, no reference counterpart).
Collected last on purpose."""
import numpy as np
import pytest

from gcpkg import gc_b200
from helpers import small_interphase, stages

pytestmark = pytest.mark.gpu


def test_noalloc_rg2_equals_the_allocating_entry_point_twice_in_a_row():
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    g = gc_b200.System(d.n, 2, 0)
    stages.setup_interphase(g, d, cfg)
    rng = np.random.default_rng(3)
    g.set_positions(x, 0)
    g.set_positions(x + rng.normal(size=x.shape) * 0.01, 1)
    # first call: every chain; second call: FEWER and different ranges, so stale scratch
    # contents from the first call would show
    r1 = d.chains
    r2 = np.array([[0, 5], [3, d.n], [7, 7]], dtype=np.int32)
    for ranges, replica in ((r1, 0), (r2, 1)):
        want = g.radius_of_gyration2(ranges, replica)
        have = g.radius_of_gyration2_noalloc(ranges, replica)
        np.testing.assert_array_equal(have, want)           # same kernel, same inputs: bit-equal
    assert (g.radius_of_gyration2_noalloc(r1, 0) > 0).all()
    # beyond the static maximum: a clear error, nothing launched
    big = np.tile(np.array([[0, 1]], dtype=np.int32), (4097, 1))
    with pytest.raises(gc_b200.api.GcError, match="exceed the static scratch"):
        g.radius_of_gyration2_noalloc(big, 0)
    assert len(g.radius_of_gyration2_noalloc(big[:4096], 0)) == 4096
    g.close()
