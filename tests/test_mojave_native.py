"""The native replay of numpy's SeedSequence / PCG64 / Generator distributions (csrc/mojave_pack.cu)
against numpy itself, and the native MOJAVE recipes against the Python packer (host only: no GPU)."""

import ctypes as C

import numpy as np
import pytest


def _words(v: int):
    out = []
    while True:
        out.append(v & 0xFFFFFFFF)
        v >>= 32
        if v == 0:
            return np.array(out, dtype=np.uint32)


def _spawn_states(entropy, spawn_key, first, n):
    from radiosim_b200 import _lib as L

    e = _words(entropy)
    k = np.concatenate([_words(x) for x in spawn_key]) if spawn_key else np.zeros(0, np.uint32)
    out = np.zeros((n, 4), dtype=np.uint64)
    L.check(L.lib().rsb_pcg64_spawn_states(e.ctypes.data, len(e), k.ctypes.data if len(k) else None, len(k),
                                           first, n, out.ctypes.data))
    return out


def _numpy_state(g):
    s = g.bit_generator.state["state"]
    m = (1 << 64) - 1
    return np.array([s["state"] >> 64, s["state"] & m, s["inc"] >> 64, s["inc"] & m], dtype=np.uint64)


@pytest.mark.parametrize("entropy,key,first", [(1337, (), 0), (0, (), 3), (2**40 + 17, (5,), 10),
                                               (123456789012345678901234567890, (2, 7), 0)])
def test_spawned_children_match_numpy(entropy, key, first):
    parent = np.random.Generator(np.random.PCG64(np.random.SeedSequence(entropy, spawn_key=key,
                                                                        n_children_spawned=first)))
    kids = parent.spawn(6)
    got = _spawn_states(entropy, key, first, 6)
    for i, k in enumerate(kids):
        assert np.array_equal(got[i], _numpy_state(k)), i


@pytest.mark.parametrize("kind,a,b,fn", [
    (0, 0, 0, lambda g, n: g.random(n)),
    (1, 0, 0, lambda g, n: g.standard_normal(n)),
    (2, 0, 0, lambda g, n: g.standard_exponential(n)),
    (3, 0.7, 0, lambda g, n: g.power(0.7, n)),
    (3, 0.5, 0, lambda g, n: g.power(0.5, n)),
    (4, 1, 4, lambda g, n: np.array([g.integers(1, 4) for _ in range(n)], dtype=np.float64)),
    (4, 0, 2, lambda g, n: np.array([g.choice([10.0, 20.0]) / 10 - 1 for _ in range(n)])),
    (5, -5.0, 5.0, lambda g, n: np.array([g.uniform(-5, 5) for _ in range(n)])),
])
def test_generator_draws_match_numpy(kind, a, b, fn):
    """Bit-identical streams, tails of the ziggurats included (2e5 draws reach every branch)."""
    from radiosim_b200 import _lib as L

    n = 200_000 if kind in (0, 1, 2, 3) else 5_000
    g = np.random.default_rng(99)
    state = _numpy_state(g).copy()
    ref = fn(g, n)
    out = np.zeros(n)
    L.check(L.lib().rsb_pcg64_draws(state.ctypes.data, kind, float(a), float(b), n, out.ctypes.data))
    assert np.array_equal(out, ref)


def test_native_recipes_match_python_packer():
    """rsb_mojave_pack against radiosim_b200.mojave's Python recipes (the restatement pinned to the
    reference through tests/golden/mojave_*): every drawn number bit-equal on 3 x 1500 children, and
    the parent generator left exactly as three spawn() calls leave it."""
    from radiosim_b200.mojave import mojave as mj

    size = 256
    amount = np.array([1500, 1500, 1500])
    parent, ref_parent = np.random.default_rng(4242), np.random.default_rng(4242)
    parent.spawn(5), ref_parent.spawn(5)                   # (children do not start at index 0)
    rec, bp = mj.native_recipes(parent, amount, size)
    kids = [ref_parent.spawn(int(a)) for a in amount]
    assert parent.bit_generator.seed_seq.n_children_spawned == ref_parent.bit_generator.seed_seq.n_children_spawned
    assert parent.bit_generator.state == ref_parent.bit_generator.state
    assert _numpy_state(parent.spawn(1)[0]).tolist() == _numpy_state(ref_parent.spawn(1)[0]).tolist()
    fns = (mj._recipe_compact, mj._recipe_one_jet, mj._recipe_two_jet)
    k = nshock = 0
    for kind, ch in enumerate(kids):
        for child in ch:
            src = fns[kind](child, size)
            r = rec[k]
            for q in range(3):
                want = tuple(src.gauss[q]) if q < len(src.gauss) else (0.0,) * 6
                assert tuple(float(v) for v in r["g"][q]) == tuple(float(v) for v in want), (kind, k, q)
            assert r["angle"] == src.angle and r["amp"] == src.amp
            assert r["flip1"] == int(src.flip1) and r["compact"] == int(src.compact)
            for q, jet in enumerate(src.jets):
                assert r["nblob"][q] == len(jet.blobs), (kind, k, q)
                assert r["strength"][q] == jet.strength and r["radius"][q] == jet.radius
                for i, b in enumerate(jet.blobs):
                    p = bp[r["blob0"][q] + i]
                    x, y, amp, width, length, a = b.args
                    assert (p["x"], p["y"], p["width"], p["length"], p["a"]) == (x, y, width, length, a)
                    if b.factor is None:
                        assert p["amp"] == amp and p["ref_r"] == -1
                    else:
                        assert p["amp"] == b.factor and (p["ref_r"], p["ref_c"]) == b.ref
                        nshock += 1
            for q in range(len(src.jets), 2):
                assert r["nblob"][q] == 0
            k += 1
    assert nshock > 1000


def test_native_recipes_decline_foreign_generators():
    from radiosim_b200.mojave import mojave as mj

    assert mj.native_recipes(np.random.Generator(np.random.MT19937(3)), [1, 1, 1], 64) is None
    assert mj.native_recipes(np.random.Generator(np.random.PCG64(np.random.SeedSequence([1, 2]))), [1, 1, 1], 64) is None
