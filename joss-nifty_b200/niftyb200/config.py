"""Global configuration, mirroring ``nifty8.config`` (reference ``src/config.py:3-40``).

Only key: ``hartley_convention`` in {"non_canonical_hartley" (default), "canonical_hartley"}.
"""

_config = {"hartley_convention": "non_canonical_hartley"}

_VALID = ("non_canonical_hartley", "canonical_hartley")


def update(key, value):
    """``nifty8.config.update`` (``src/config.py:8-40``): same keys, same errors."""
    if key != "hartley_convention":
        raise ValueError(f"Unknown configuration key {key!r}")
    if value not in _VALID:
        raise ValueError(f"Invalid value {value!r} for {key!r}; expected one of {_VALID}")
    _config[key] = value


def hartley_convention():
    return _config["hartley_convention"]


def convention_code(name=None):
    """0 = non-canonical (Re F + Im F), 1 = canonical (Re F - Im F)."""
    name = hartley_convention() if name is None else name
    if name not in _VALID:
        raise ValueError(f"invalid hartley convention {name!r}")
    return 0 if name == "non_canonical_hartley" else 1
