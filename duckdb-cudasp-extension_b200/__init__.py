"""cudasp-b200: B200-native cudasp_scan hot path.  csrc/ = CUDA kernels + C ABI, host/ = Python mirror
of the reference's table-function interface above that ABI, duckdb_ext/ = the C++ DuckDB shim."""
from .host import *  # noqa: F401,F403
