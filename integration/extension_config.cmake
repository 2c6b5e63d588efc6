# DuckDB extension config: the `rpt` extension with the GPU bloom-filter body (integration/CMakeLists.txt)
duckdb_extension_load(rpt
    SOURCE_DIR ${CMAKE_CURRENT_LIST_DIR}
    DONT_LINK
)
