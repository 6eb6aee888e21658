"""Minimal obspy stand-in for running the reference in the build container (test infrastructure only)."""
from .core.stream import Stream, Trace, Stats, UTCDateTime  # noqa: F401
