# Included by DuckDB's build (DUCKDB_EXTENSION_CONFIGS): the REFERENCE's `cudasp` extension, built by oracle/ref_ext/CMakeLists.txt.
duckdb_extension_load(cudasp
    SOURCE_DIR ${CMAKE_CURRENT_LIST_DIR}
    INCLUDE_DIR /root/reference/src/include
    EXTENSION_VERSION reference-98601999
)
