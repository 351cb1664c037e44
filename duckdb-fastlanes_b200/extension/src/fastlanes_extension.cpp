// fastlanes_extension.cpp — loads the B200 FastLanes extension into DuckDB: the read_fls table function, the `fls` copy
// format and the extension's settings.  Counterpart of the reference's src/fastlanes_extension.cpp:10-36 (same entry
// symbol fastlanes_duckdb_cpp_init, same extension name).
#define DUCKDB_EXTENSION_MAIN

#include "fastlanes_extension.hpp"

#include "duckdb/main/config.hpp"
#include "duckdb/main/extension/extension_loader.hpp"
#include "flsgpu.h"

namespace duckdb {

static void RegisterSettings(ExtensionLoader &loader) {
	auto &config = DBConfig::GetConfig(loader.GetDatabaseInstance());
	// a worker thread is handed this many row groups at once: one pipelined H2D -> decode -> D2H call per batch
	config.AddExtensionOption("fastlanes_rowgroups_per_batch",
	                          "Row groups a read_fls worker decodes per GPU call (the reference hands out one at a time)",
	                          LogicalType::UBIGINT, Value::UBIGINT(32));
	config.AddExtensionOption("fastlanes_compressed_vectors",
	                          "Hand vectors the file marks CONSTANT or arithmetic SEQUENCE to DuckDB as CONSTANT / SEQUENCE vectors "
	                          "instead of decoding and copying them",
	                          LogicalType::BOOLEAN, Value::BOOLEAN(true));
	config.AddExtensionOption("fastlanes_select_on_device",
	                          "Apply the filters the device evaluates on the device: only the rows that pass are copied to the host "
	                          "(off: every row of the projected columns comes back with a selection mask)",
	                          LogicalType::BOOLEAN, Value::BOOLEAN(true));
	config.AddExtensionOption("fastlanes_table_cache_mb",
	                          "Opened .fls files stay open - container validated, image page-locked, compressed chunks resident in HBM - "
	                          "for the next query until the cached files exceed this many MiB (0: no caching)",
	                          LogicalType::UBIGINT, Value::UBIGINT(65536));
	config.AddExtensionOption("fastlanes_gpu_devices",
	                          "Comma separated CUDA device ordinals read_fls shards row groups over (empty: device 0)",
	                          LogicalType::VARCHAR, Value(""));
}

static void LoadFastLanes(ExtensionLoader &loader) {
	RegisterSettings(loader);
	FastLanesScanFunction::Register(loader);
	FastLanesCopyFunction::Register(loader);
	FastLanesAggregatePushdown::Register(loader);
}

void FastlanesExtension::Load(ExtensionLoader &loader) {
	LoadFastLanes(loader);
}

std::string FastlanesExtension::Name() {
	return "fastlanes";
}

std::string FastlanesExtension::Version() const {
	return flsgpu_version();
}

} // namespace duckdb

extern "C" {

DUCKDB_CPP_EXTENSION_ENTRY(fastlanes, loader) {
	duckdb::LoadFastLanes(loader);
}
}
