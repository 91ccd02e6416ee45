# Included by DuckDB's build (DUCKDB_EXTENSION_CONFIGS): build the B200-native `cudasp` extension from this directory.
duckdb_extension_load(cudasp
    SOURCE_DIR ${CMAKE_CURRENT_LIST_DIR}
    INCLUDE_DIR ${CMAKE_CURRENT_LIST_DIR}
    EXTENSION_VERSION b200-0.1
)
