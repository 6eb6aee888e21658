"""Empty stand-in for matplotlib.pyplot (plotting is out of scope)."""


def figure(*a, **k):  # pragma: no cover - never called on the hot path
    raise RuntimeError("plotting is unavailable in the oracle stub")
