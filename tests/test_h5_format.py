"""The ``.h5`` writer against the reference's h5py-written fixture and the format specification.

No HDF5 library exists in the build container or on the GPU box (``h5py``, ``tables``,
``netCDF4``, ``libhdf5``: all absent, probed on both), so two external anchors are used instead:

* the reference's own fixture ``tests/ppdisks_128.h5`` (written by h5py): writing its dataset
  back with ``write_h5`` must give the same file **byte for byte**;
* ``tests/h5spec.py``, a field-by-field walker of the HDF5 format specification that shares no
  code with the writer or its reader, run over a shard of every mode.
"""

import os

import numpy as np
import pytest

from radiosim_b200.h5min import read_h5, write_h5

import h5spec

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURE = os.path.join(HERE, "golden", "ppdisks_128.h5")


def test_rewriting_the_reference_fixture_is_byte_identical(tmp_path):
    y = read_h5(FIXTURE)["y"]
    out = tmp_path / "again.h5"
    write_h5(out, {"y": y})
    assert out.read_bytes() == open(FIXTURE, "rb").read()


def test_reference_fixture_passes_the_spec_walker():
    w = h5spec.walk(FIXTURE)
    (d,) = w["datasets"]
    assert d["name"] == "y" and d["shape"] == (9, 128, 128)
    assert d["dtype"]["class"] == 1 and d["dtype"]["size"] == 4
    assert d["message_types"] == [0x01, 0x03, 0x05, 0x08, 0x00]   # what h5py itself wrote
    assert d["data"] == (0x800, 9 * 128 * 128 * 4)


def _shards(tmp_path):
    rng = np.random.default_rng(0)
    return {
        # save_sky_distribution_bundle, reference utils/utils.py:344-364
        "sky": {"x": rng.random((3, 1, 16, 16)), "y": rng.random((3, 19, 8))},
        # _save_mojave_bundle, utils/utils.py:367-383 (date: h5py.string_dtype('utf-8', 10))
        "mojave": {"galaxies": rng.random((4, 8, 8)), "galaxy_classes": np.arange(4, dtype=np.int64),
                   "RA": rng.random(4), "DEC": rng.random(4),
                   "date": np.array(["1999-01-02", "2004-10-30", "2011-03-04", "2020-12-31"], dtype="S10")},
        # simulations/ppdisks.py:84-89 (y float32, metadata / params: scalar variable-length strings)
        "ppdisk": {"y": rng.random((2, 8, 8)).astype(np.float32),
                   "metadata": str([dict(index=0, cell_size=0.003, src_ra=1.0, src_dec=-2.0)]),
                   "params": str(dict(seed=42, N=2, img_size=8)) + " °"},
    }


@pytest.mark.parametrize("mode", ["sky", "mojave", "ppdisk"])
def test_written_shards_pass_the_spec_walker(tmp_path, mode):
    data = _shards(tmp_path)[mode]
    path = tmp_path / f"{mode}.h5"
    write_h5(path, data)
    w = h5spec.walk(path)
    assert [d["name"] for d in w["datasets"]] == sorted(data)
    back = read_h5(path)
    for d in w["datasets"]:
        v = data[d["name"]]
        if isinstance(v, str):
            assert d["shape"] == () and d["dtype"]["class"] == 9 and d["dtype"]["cset"] == 1
            assert d["strings"] == [v] and back[d["name"]] == v
        else:
            assert d["shape"] == v.shape
            a, n = d["data"]
            raw = open(path, "rb").read()[a : a + n]
            assert raw == np.ascontiguousarray(v).tobytes()
            assert np.array_equal(back[d["name"]], v)
            cls = {"f": 1, "i": 0, "S": 3}[v.dtype.kind]
            assert d["dtype"]["class"] == cls and d["dtype"]["size"] == v.dtype.itemsize
            if cls == 3:
                assert (d["dtype"]["strpad"], d["dtype"]["cset"]) == (1, 1)   # null-padded UTF-8
            if cls == 0:
                assert d["dtype"]["signed"]
        # same message list, flags and fill-value settings as the dataset h5py wrote
        ref = h5spec.walk(FIXTURE)["datasets"][0]
        assert h5spec.signature(d)[3:6] == h5spec.signature(ref)[3:6]


def test_spec_walker_catches_damage(tmp_path):
    """The walker is not a rubber stamp: single-field corruptions of a good file are reported."""
    good = bytearray(open(FIXTURE, "rb").read())
    h5spec.walk(bytes(good))
    for off, val in [(0x28, 0x01),     # end-of-file address
                     (0x324, 0x01),    # message count of the dataset header
                     (0x378, 0x21),    # datatype version
                     (0x3a9, 0x02),    # layout class
                     (0x2b8, 0x18),    # local-heap free-list head
                     (0x436, 0x03)]:   # symbol count
        bad = bytearray(good)
        bad[off] ^= val
        with pytest.raises(AssertionError):
            h5spec.walk(bytes(bad))
