"""minicfd_b200: B200-native DG/FR residual + RK hot path behind miniCFD's API."""
from . import mesh  # noqa: F401
