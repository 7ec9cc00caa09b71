"""CPU tests of the host-side jet packer: draw-for-draw alignment with the oracle."""

import numpy as np
import pytest

from oracle import jet as ojet
from radiosim_b200.jets.packer import pack_jets, source_lists


@pytest.mark.parametrize("size,scaling,seed", [(128, "normalize", 1337), (64, "normalize", 3),
                                               (200, "mojave", 5), (256, "normalize", 11)])
def test_packer_matches_oracle_stream(size, scaling, seed):
    n = 24
    np.random.seed(seed)
    trace = []
    _, y = ojet.create_jet((n, 3, size, size), (3, 10), scaling, "list", trace=trace)
    end_oracle = np.random.rand()

    np.random.seed(seed)
    b = pack_jets(n, size, (3, 10), scaling)
    end_packer = np.random.rand()
    assert end_packer == end_oracle  # same number of draws, same order

    zoomed = np.array([t["zoomed"] for t in trace])
    assert np.array_equal(zoomed, b.zoom_pad >= 0)
    assert np.array_equal(np.array([t["pad"] for t in trace])[zoomed], b.zoom_pad[zoomed])
    if scaling == "normalize":
        assert zoomed.any() and (~zoomed).any()

    # with the oracle's margin and peak the list targets are bit-identical
    margin = np.array([t["margin"] for t in trace])
    peak = np.array([t["jet_max"] for t in trace])
    out = source_lists(b, margin, peak, scaling == "normalize")
    assert np.array_equal(out, y)


def test_packer_slot_layout():
    np.random.seed(0)
    b = pack_jets(5, 128, (3, 6))
    assert b.params.shape == (5, 11, 6) and b.zoom_pad.dtype == np.int32
    assert np.all(b.params[:, 0, 0] == 1.0)  # core: amp 1, beta 0 -> boost 1
