"""nbls-b200: the narrow-band least-squares (NBLS) array-processing hot path on NVIDIA B200.

Public surface (same names and argument meaning as amiezzi/narrow_band_least_squares):
    narrow_band_least_squares, narrow_band_least_squares_parallel   (narrow_band_least_squares.py)
    ltsva                                                           (lts_array.ltsva)
    narrow_band_least_squares_batch                                 (the driver over a list of independent arrays)
    helpers: get_freqlist, get_winlenlist, filter_data, make_float, write_txtfile, read_txtfile, get_rij
Everything numerical runs in hand-written sm_100a CUDA kernels behind the C ABI of include/nbls.h
(libnbls.so, built in-tree under _lib/); importing this package does not need a GPU, calling it does.
"""
from .helpers import (filter_data, get_freqlist, get_rij, get_winlenlist, make_float, read_txtfile,  # noqa: F401
                      write_txtfile)
from .batch import narrow_band_least_squares_batch  # noqa: F401
from .ltsva import ltsva  # noqa: F401
from .narrow_band_least_squares import narrow_band_least_squares, narrow_band_least_squares_parallel  # noqa: F401

__version__ = "0.2.0"
