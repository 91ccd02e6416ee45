from .binding import CudaspError, Query, lib, test_field, debug_gtab, LIB_PATH, EXPORTS, profile_stages, STAGES, KEYS_XY64, KEYS_SEC1_33, MatchDetail  # noqa: F401
from .table_function import (CudaspScanBind, InvalidInputException, IOException, cudasp_scan)  # noqa: F401
from .arrow import (ARROW_EXPORTS, cudasp_scan_arrow, scan_columns, scan_record_batch, scan_parquet, scan_stream, scan_table)  # noqa: F401
from .parquet import PARQUET_EXPORTS, scan_parquet_files  # noqa: F401
