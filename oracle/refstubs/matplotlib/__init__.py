"""Empty stand-in: the reference imports matplotlib.pyplot at module top but the hot path never plots."""
