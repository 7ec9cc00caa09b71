"""TOML configuration with radiosim's key names (reference ``src/radiosim/io/config.py``).

Only the keys the hot path reads are interpreted; unknown keys are ignored like in the
reference.  As there (``Field(default_value=...)`` is not a default in pydantic), the dataset
sizes are REQUIRED.  ``"none"`` / ``false`` mean ``None`` for ``seed`` and ``threads``.
"""

from __future__ import annotations

import tomllib
from pathlib import Path
from types import SimpleNamespace

__all__ = ["Config", "ConfigError"]


class ConfigError(ValueError):
    pass


_DEFAULTS = {
    "general": dict(verbose=False, seed=None, threads=None, device="cuda"),
    "paths": dict(outpath="./build/example_data/"),
    "dataset": dict(noise=True, noise_level=[0.0, 15.0]),
    "jet": dict(training_type="list", num_jet_components=[3, 10], scaling="normalize"),
    "survey": dict(num_sources=20, class_distribution=[2, 1, 2], scale_sources=True),
    "mojave": dict(class_ratio=[1, 1, 1]),
}
_REQUIRED = {"dataset": ["bundles_train", "bundles_valid", "bundles_test", "bundle_size", "img_size"]}
_LENGTHS = {("jet", "num_jet_components"): 2, ("survey", "class_distribution"): 3,
            ("mojave", "class_ratio"): 3, ("dataset", "noise_level"): 2}
_CHOICES = {("jet", "training_type"): ("list", "gauss", "clean"), ("jet", "scaling"): ("normalize", "mojave")}


class Config(SimpleNamespace):
    @classmethod
    def from_dict(cls, data: dict) -> "Config":
        out = {}
        for section, defaults in _DEFAULTS.items():
            merged = dict(defaults)
            merged.update(data.get(section, {}) or {})
            for key in _REQUIRED.get(section, []):
                if key not in merged:
                    raise ConfigError(f"[{section}] {key}: field required")
            out[section] = merged
        for (section, key), n in _LENGTHS.items():
            if len(out[section][key]) != n:
                raise ConfigError(f"Expected '{key}' to be of length {n}!")
        for (section, key), allowed in _CHOICES.items():
            if out[section][key] not in allowed:
                raise ConfigError(f"[{section}] {key} must be one of {allowed}")
        g = out["general"]
        for key in ("seed", "threads"):
            if g[key] in ("none", False):
                g[key] = None
        d = out["dataset"]
        if d["img_size"] < 64 or d["bundle_size"] < 1 or min(d["bundles_train"], d["bundles_valid"], d["bundles_test"]) < 0:
            raise ConfigError("[dataset] img_size >= 64, bundle_size >= 1, bundles_* >= 0 required")
        if out["paths"]["outpath"] in (None, False, "none", ""):
            raise ConfigError("'outpath' cannot be empty!")
        out["paths"]["outpath"] = Path(out["paths"]["outpath"])
        pp = data.get("ppdisk", data.get("ppdisks", {})) or {}
        out["ppdisk"] = dict(
            metadata=SimpleNamespace(**{**dict(alpha_range=[0.0, 180.0], ratio_range=[3.0, 15.0],
                                               size_ratio_range=[0.1, 1.0]), **pp.get("metadata", {})}),
            dataset=SimpleNamespace(**{**dict(file_prefix="ppdisks"), **pp.get("dataset", {})}))
        return cls(**{k: SimpleNamespace(**v) for k, v in out.items()})

    @classmethod
    def from_toml(cls, path) -> "Config":
        with open(path, "rb") as f:
            return cls.from_dict(tomllib.load(f))
