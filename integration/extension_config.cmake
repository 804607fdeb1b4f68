# Included by DuckDB's build (-DDUCKDB_EXTENSION_CONFIGS=...): the vss extension over libgvss (see CMakeLists.txt here).
set(VSS_REFERENCE_DIR "/root/reference" CACHE PATH "checkout of duckdb/duckdb_vss")
duckdb_extension_load(vss
        SOURCE_DIR ${CMAKE_CURRENT_LIST_DIR}
        INCLUDE_DIR ${VSS_REFERENCE_DIR}/src/include
        )
