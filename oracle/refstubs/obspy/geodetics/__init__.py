from .base import calc_vincenty_inverse  # noqa: F401
