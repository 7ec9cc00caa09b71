"""Command line: ``radiosim simulate CONFIG --mode {mojave,jet,survey,ppdisk}`` plus the legacy
spellings ``radiosim-simulate`` / ``radiosim-ppdisk`` (reference ``src/radiosim/tools``,
``README.rst:69,90``).  Under ``torchrun`` every rank renders and writes its own shards."""

from __future__ import annotations

import argparse
import sys
import tomllib


def simulate(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="radiosim simulate")
    ap.add_argument("configuration_path")
    ap.add_argument("--mode", default="mojave", choices=["mojave", "jet", "survey", "ppdisk"])
    args = ap.parse_args(argv)
    from .. import _lib as L
    from ..sharder import rank_info

    _, world, local = rank_info()
    if world > 1:  # one process per GPU: keep each one's host buffers on its GPU's socket
        from ..sharder import bind_near_gpu

        bind_near_gpu(local)
    L.require_device()
    L.set_device(local)
    if args.mode == "ppdisk":
        from ..simulations import simulate_ppdisks

        with open(args.configuration_path, "rb") as f:
            written = simulate_ppdisks(tomllib.load(f))
    else:
        from ..io import Config
        from ..simulations import simulate_sky_distributions

        written = simulate_sky_distributions(Config.from_toml(args.configuration_path), args.mode)
    print(f"wrote {len(written)} shard(s)")
    return 0


def ppdisk(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    return simulate(argv + ["--mode", "ppdisk"])


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if argv and argv[0] == "simulate":
        return simulate(argv[1:])
    print("usage: radiosim simulate CONFIG --mode {mojave,jet,survey,ppdisk}", file=sys.stderr)
    return 2


if __name__ == "__main__":
    sys.exit(main())
