// fls_scan_filters.cpp — pushed-down TableFilters of a read_fls scan: what goes to the device, what stays on the host.
//
// The reference evaluates every pushed-down filter on the worker's CPU after decode (FilterExecutor::Apply,
// src/reader/filter_executor.cpp:19-60) and prunes row groups from min/max statistics for constant comparisons
// (RowGroupFilter, src/reader/row_group_filter.cpp:96-199).  Here constant comparisons on physical columns — the filters
// zone maps can prune and the fused decode+compare kernels evaluate (flsgpu_filter_mask) — go to the device as
// flsgpu_filter entries; everything else is kept as a host filter and runs through the same DuckDB call the reference uses.
#include "fls_file_reader.hpp"

#include "duckdb/planner/filter/conjunction_filter.hpp"
#include "duckdb/planner/filter/constant_filter.hpp"
#include "duckdb/planner/filter/optional_filter.hpp"

#include <cstring>

namespace duckdb {

FastLanesHostFilter::FastLanesHostFilter(ClientContext &context, idx_t chunk_column_p, TableFilter &filter_p)
    : chunk_column(chunk_column_p), filter(filter_p), filter_state(TableFilterState::Initialize(context, filter_p)) {
}

static bool ComparisonToDevice(ExpressionType type, int32_t &op) {
	switch (type) {
	case ExpressionType::COMPARE_EQUAL:
		op = FLSGPU_CMP_EQ;
		return true;
	case ExpressionType::COMPARE_NOTEQUAL:
		op = FLSGPU_CMP_NE;
		return true;
	case ExpressionType::COMPARE_LESSTHAN:
		op = FLSGPU_CMP_LT;
		return true;
	case ExpressionType::COMPARE_LESSTHANOREQUALTO:
		op = FLSGPU_CMP_LE;
		return true;
	case ExpressionType::COMPARE_GREATERTHAN:
		op = FLSGPU_CMP_GT;
		return true;
	case ExpressionType::COMPARE_GREATERTHANOREQUALTO:
		op = FLSGPU_CMP_GE;
		return true;
	default:
		return false;
	}
}

// The constant as the physical word of the column (flsgpu_filter.constant): integers sign- / zero-extended to 64 bits,
// FLOAT bits in the low word, DOUBLE bits.  False when the constant is not of the column's type or does not fit the
// column's stored width (such a comparison is left to the host).
static bool ConstantToPhysical(const Value &constant, const fls_column_desc &column, const LogicalType &type, uint64_t &word) {
	if (constant.IsNull() || constant.type() != type) {
		return false;
	}
	switch (type.InternalType()) {
	case PhysicalType::BOOL:
		word = BooleanValue::Get(constant) ? 1 : 0;
		return true;
	case PhysicalType::INT8:
		word = static_cast<uint64_t>(static_cast<int64_t>(constant.GetValueUnsafe<int8_t>()));
		return true;
	case PhysicalType::INT16:
		word = static_cast<uint64_t>(static_cast<int64_t>(constant.GetValueUnsafe<int16_t>()));
		return true;
	case PhysicalType::INT32:
		word = static_cast<uint64_t>(static_cast<int64_t>(constant.GetValueUnsafe<int32_t>()));
		return true;
	case PhysicalType::INT64:
		word = static_cast<uint64_t>(constant.GetValueUnsafe<int64_t>());
		return true;
	case PhysicalType::UINT64: {
		// UINT8..UINT64 all read back as UBIGINT (translation_utils.cpp:19-24); the device compares at the stored width
		word = constant.GetValueUnsafe<uint64_t>();
		return column.phys_bits >= 64 || word < (uint64_t(1) << column.phys_bits);
	}
	case PhysicalType::FLOAT: {
		const float f = constant.GetValueUnsafe<float>();
		uint32_t bits;
		std::memcpy(&bits, &f, sizeof(bits));
		word = bits;
		return true;
	}
	case PhysicalType::DOUBLE: {
		const double d = constant.GetValueUnsafe<double>();
		std::memcpy(&word, &d, sizeof(word));
		return true;
	}
	default:
		return false;
	}
}

bool FastLanesFilterPushdown::ToDeviceFilter(const TableFilter &filter, const fls_column_desc &column, const LogicalType &type,
                                             uint32_t col, vector<flsgpu_filter> &out, vector<string> &strings) {
	switch (filter.filter_type) {
	case TableFilterType::CONSTANT_COMPARISON: {
		auto &constant_filter = filter.Cast<ConstantFilter>();
		flsgpu_filter f {};
		f.col = col;
		if (!ComparisonToDevice(constant_filter.comparison_type, f.op)) {
			return false;
		}
		if (type.InternalType() == PhysicalType::VARCHAR) {
			if (constant_filter.constant.IsNull() || constant_filter.constant.type().InternalType() != PhysicalType::VARCHAR) {
				return false;
			}
			strings.push_back(StringValue::Get(constant_filter.constant)); // bytes outlive the scan in the local state
			f.str_len = strings.back().size();
			f.constant = strings.size() - 1; // index for now; pointers are fixed up once the vector stops growing
			out.push_back(f);
			return true;
		}
		if (!ConstantToPhysical(constant_filter.constant, column, type, f.constant)) {
			return false;
		}
		out.push_back(f);
		return true;
	}
	case TableFilterType::CONJUNCTION_AND: {
		auto &conjunction = filter.Cast<ConjunctionAndFilter>();
		const auto n_out = out.size(), n_strings = strings.size();
		for (auto &child : conjunction.child_filters) {
			if (!ToDeviceFilter(*child, column, type, col, out, strings)) {
				out.resize(n_out);
				strings.resize(n_strings);
				return false;
			}
		}
		return true;
	}
	case TableFilterType::IS_NOT_NULL:
		return true; // FastLanes files hold no NULLs (fls_reader.cpp:199,208): always true
	default:
		return false;
	}
}

void FastLanesFilterPushdown::Split(ClientContext &context, BaseFileReader &reader, const vector<fls_column_desc> &column_descs,
                                    idx_t n_physical_columns, FastLanesReadLocalState &ls) {
	for (auto &entry : reader.filters->filters) {
		const idx_t chunk_column = entry.first; // position in column_ids = column of the scan chunk (filter_executor.cpp:42)
		auto &filter = *entry.second;
		if (chunk_column >= reader.column_ids.size()) {
			throw InternalException("read_fls: filter on a column that is not part of the scan");
		}
		const auto local_id = reader.column_ids[MultiFileLocalIndex(chunk_column)].GetId();
		const bool physical = local_id < n_physical_columns;
		const bool converted = reader.expression_map.find(local_id) != reader.expression_map.end();
		if (physical && !converted) {
			const auto n_strings = ls.device_filter_strings.size();
			if (ToDeviceFilter(filter, column_descs[local_id], reader.columns[local_id].type, NumericCast<uint32_t>(local_id),
			                   ls.device_filters, ls.device_filter_strings)) {
				continue;
			}
			ls.device_filter_strings.resize(n_strings);
			// an optional filter (e.g. the probe-side min/max of a join) helps pruning when the device can take it and is
			// skipped otherwise: executing it is not required for correctness
			if (filter.filter_type == TableFilterType::OPTIONAL_FILTER) {
				auto &optional = filter.Cast<OptionalFilter>();
				if (optional.child_filter &&
				    !ToDeviceFilter(*optional.child_filter, column_descs[local_id], reader.columns[local_id].type,
				                    NumericCast<uint32_t>(local_id), ls.device_filters, ls.device_filter_strings)) {
					ls.device_filter_strings.resize(n_strings);
				}
				continue;
			}
		}
		ls.host_filters.emplace_back(context, chunk_column, filter);
	}
	// string constants: the entries point at the local state's copies
	for (auto &f : ls.device_filters) {
		if (column_descs[f.col].phys_bits == 128) {
			auto &bytes = ls.device_filter_strings[f.constant];
			f.str_data = reinterpret_cast<const uint8_t *>(bytes.data());
			f.str_len = bytes.size();
			f.constant = 0;
		}
	}
}

} // namespace duckdb
