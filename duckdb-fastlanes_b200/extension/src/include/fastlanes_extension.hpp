// fastlanes_extension.hpp — entry class DuckDB's generated loader instantiates (LoadStaticExtension<FastlanesExtension>)
// for the B200 FastLanes extension.  Same extension name, functions and options as the reference's
// src/include/fastlanes_extension.hpp / src/fastlanes_extension.cpp:10-36; the decode work behind read_fls runs in
// libflsgpu.so (include/flsgpu.h).
#pragma once

#include "duckdb.hpp"
#include "duckdb/main/extension.hpp"

namespace duckdb {

class FastlanesExtension : public Extension {
public:
	void Load(ExtensionLoader &loader) override;
	std::string Name() override;
	std::string Version() const override;
};

//! read_fls (reference src/read_fls.cpp:32-46)
struct FastLanesScanFunction {
	static void Register(ExtensionLoader &loader);
};
//! Ungrouped SUM / COUNT / MIN / MAX over read_fls answered by fused decode + aggregate kernels (no reference counterpart)
struct FastLanesAggregatePushdown {
	static void Register(ExtensionLoader &loader);
};
//! COPY ... (FORMAT fls) (reference src/write_fls.cpp:8-29)
struct FastLanesCopyFunction {
	static void Register(ExtensionLoader &loader);
};

} // namespace duckdb
