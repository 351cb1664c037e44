# Included by DuckDB's build (-DDUCKDB_EXTENSION_CONFIGS=<this file>): links the B200 `fastlanes` extension into the
# duckdb shell and the unittest runner and registers its SQLLogicTests.  Stands where the reference's
# extension_config.cmake:4-7 stands.
duckdb_extension_load(fastlanes
        SOURCE_DIR ${CMAKE_CURRENT_LIST_DIR}
        INCLUDE_DIR ${CMAKE_CURRENT_LIST_DIR}/src/include
        LOAD_TESTS
)
# dbgen for BASELINE configs 3/5 (TPC-H lineitem written through COPY ... (FORMAT fls))
duckdb_extension_load(tpch)
