// fls_aggregate_pushdown.cpp — SELECT SUM(c), COUNT(*), MIN(c), MAX(c) ... FROM read_fls('file') [WHERE constant comparisons]
// answered by the fused decode + aggregate kernels (flsgpu_aggregate_to_host) instead of a scan that materialises every
// value: an OptimizerExtension (duckdb/optimizer/optimizer_extension.hpp:31-43) replaces
//     AGGREGATE (no groups) -> [PROJECTION of column references] -> GET read_fls
// by a one-row table function.  No counterpart in the reference: there every value is decoded into DuckDB vectors
// (fls_reader.cpp:436-472) and DuckDB's ungrouped aggregate consumes them.  Only compressed bytes cross PCIe (none at all
// when the table is already resident, FlsTableCache) and 64 bytes per column come back.
//
// The rewrite keeps DuckDB's semantics exactly or does not happen: integer / DECIMAL sums are the 128-bit sums DuckDB
// computes (sum.cpp:161-210: HUGEINT, DECIMAL(38, s)), FastLanes files hold no NULLs (COUNT(c) = COUNT(*)), MIN / MAX come
// from the chunks' zone maps and are pushed only when every chunk carries statistics and no filter applies; floating point
// sums (order dependent), DISTINCT / FILTER / ORDER BY aggregates, GROUP BY, expressions over columns, several files,
// virtual columns and filters the device cannot evaluate completely leave the plan untouched.
#include "fastlanes_extension.hpp"
#include "fls_file_reader.hpp"

#include "duckdb/common/multi_file/multi_file_states.hpp"
#include "duckdb/main/config.hpp"
#include "duckdb/optimizer/optimizer_extension.hpp"
#include "duckdb/planner/expression/bound_aggregate_expression.hpp"
#include "duckdb/planner/expression/bound_columnref_expression.hpp"
#include "duckdb/planner/operator/logical_aggregate.hpp"
#include "duckdb/planner/operator/logical_get.hpp"
#include "duckdb/planner/operator/logical_projection.hpp"

namespace duckdb {

namespace {

enum class FlsAggregateKind : uint8_t { COUNT, SUM, MIN, MAX };

struct FlsAggregateSpec {
	FlsAggregateKind kind;
	uint32_t column = 0; // physical column of the file (unused for COUNT)
	LogicalType return_type;
};

struct FlsAggregateBindData : public TableFunctionData {
	shared_ptr<FlsCachedTable> table;
	string path;
	vector<FlsAggregateSpec> aggregates;
	vector<flsgpu_filter> filters;
	vector<string> filter_strings;
	uint32_t n_rowgroups = 0;
	uint64_t n_tuples = 0;
	//! a numeric column for COUNT(*) WHERE ... when no column is aggregated (the count of rows that pass rides on it)
	optional_idx count_carrier;
};

struct FlsAggregateGlobalState : public GlobalTableFunctionState {
	bool done = false;
	idx_t MaxThreads() const override {
		return 1;
	}
};

unique_ptr<GlobalTableFunctionState> FlsAggregateInit(ClientContext &, TableFunctionInitInput &) {
	return make_uniq<FlsAggregateGlobalState>();
}

hugeint_t SumOf(const flsgpu_agg_result &r) {
	hugeint_t h;
	h.lower = r.sum_lo;
	h.upper = static_cast<int64_t>(r.sum_hi);
	return h;
}

Value ExtremeValue(const LogicalType &type, int64_t v) {
	switch (type.id()) {
	case LogicalTypeId::UBIGINT:
		return Value::UBIGINT(static_cast<uint64_t>(v));
	case LogicalTypeId::DATE:
		return Value::DATE(date_t(NumericCast<int32_t>(v)));
	case LogicalTypeId::TIMESTAMP:
		return Value::TIMESTAMP(timestamp_t(v));
	case LogicalTypeId::DECIMAL: {
		const auto width = DecimalType::GetWidth(type), scale = DecimalType::GetScale(type);
		switch (type.InternalType()) {
		case PhysicalType::INT16:
			return Value::DECIMAL(NumericCast<int16_t>(v), width, scale);
		case PhysicalType::INT32:
			return Value::DECIMAL(NumericCast<int32_t>(v), width, scale);
		default:
			return Value::DECIMAL(v, width, scale);
		}
	}
	default:
		return Value::Numeric(type, v);
	}
}

void FlsAggregateScan(ClientContext &, TableFunctionInput &data, DataChunk &output) {
	auto &bind = data.bind_data->Cast<FlsAggregateBindData>();
	auto &state = data.global_state->Cast<FlsAggregateGlobalState>();
	if (state.done) {
		output.SetCardinality(0);
		return;
	}
	state.done = true;
	// one fused launch set per distinct column; COUNT(*) rides on any of them
	vector<uint32_t> columns;
	vector<idx_t> result_of(bind.aggregates.size(), 0);
	for (idx_t i = 0; i < bind.aggregates.size(); i++) {
		auto &spec = bind.aggregates[i];
		if (spec.kind == FlsAggregateKind::COUNT) {
			continue;
		}
		idx_t at = 0;
		for (; at < columns.size() && columns[at] != spec.column; at++) {
		}
		if (at == columns.size()) {
			columns.push_back(spec.column);
		}
		result_of[i] = at;
	}
	vector<flsgpu_agg_result> results(MaxValue<idx_t>(columns.size(), 1));
	uint64_t count = bind.n_tuples;
	const bool need_device = !columns.empty() || !bind.filters.empty();
	if (need_device && bind.n_rowgroups > 0) {
		if (columns.empty()) { // COUNT(*) WHERE ...
			columns.push_back(NumericCast<uint32_t>(bind.count_carrier.GetIndex()));
		}
		FlsCheck(flsgpu_aggregate_to_host(bind.table->handles[0], columns.data(), NumericCast<uint32_t>(columns.size()),
		                                  bind.filters.empty() ? nullptr : bind.filters.data(), NumericCast<uint32_t>(bind.filters.size()), 0,
		                                  bind.n_rowgroups, results.data()),
		         "flsgpu_aggregate_to_host");
		count = results[0].count;
	}
	output.SetCardinality(1);
	for (idx_t i = 0; i < bind.aggregates.size(); i++) {
		auto &spec = bind.aggregates[i];
		const auto &r = results[result_of[i]];
		switch (spec.kind) {
		case FlsAggregateKind::COUNT:
			output.SetValue(i, 0, Value::BIGINT(NumericCast<int64_t>(count)));
			break;
		case FlsAggregateKind::SUM:
			if (count == 0) {
				output.SetValue(i, 0, Value(spec.return_type));
			} else if (spec.return_type.id() == LogicalTypeId::DECIMAL) {
				output.SetValue(i, 0, Value::DECIMAL(SumOf(r), DecimalType::GetWidth(spec.return_type), DecimalType::GetScale(spec.return_type)));
			} else {
				output.SetValue(i, 0, Value::HUGEINT(SumOf(r)));
			}
			break;
		case FlsAggregateKind::MIN:
		case FlsAggregateKind::MAX:
			if (count == 0) {
				output.SetValue(i, 0, Value(spec.return_type));
			} else {
				output.SetValue(i, 0, ExtremeValue(spec.return_type, spec.kind == FlsAggregateKind::MIN ? r.min_val : r.max_val));
			}
			break;
		}
	}
}

InsertionOrderPreservingMap<string> FlsAggregateToString(TableFunctionToStringInput &input) {
	InsertionOrderPreservingMap<string> result;
	auto &bind = input.bind_data->Cast<FlsAggregateBindData>();
	result["File"] = bind.path;
	result["Aggregates"] = to_string(bind.aggregates.size()) + " fused on the GPU";
	if (!bind.filters.empty()) {
		result["Filters"] = to_string(bind.filters.size()) + " evaluated on the GPU";
	}
	return result;
}

bool SummableType(const LogicalType &type) {
	switch (type.id()) {
	case LogicalTypeId::TINYINT:
	case LogicalTypeId::SMALLINT:
	case LogicalTypeId::INTEGER:
	case LogicalTypeId::BIGINT:
	case LogicalTypeId::UBIGINT: // every unsigned width reads back as UBIGINT (translation_utils.cpp:19-24)
		return true;
	case LogicalTypeId::DECIMAL:
		return type.InternalType() != PhysicalType::INT128;
	default:
		return false;
	}
}

bool OrderedType(const LogicalType &type) {
	return SummableType(type) || type.id() == LogicalTypeId::DATE || type.id() == LogicalTypeId::TIMESTAMP;
}

// The plan below `op` when it is AGGREGATE -> [PROJECTION] -> GET read_fls of the supported shape; nullptr otherwise.
unique_ptr<LogicalOperator> TryRewriteAggregate(ClientContext &context, LogicalAggregate &aggr) {
	if (!aggr.groups.empty() || aggr.grouping_sets.size() > 1 || !aggr.grouping_functions.empty() || aggr.expressions.empty() ||
	    aggr.children.size() != 1) {
		return nullptr;
	}
	LogicalOperator *child = aggr.children[0].get();
	LogicalProjection *projection = nullptr;
	if (child->type == LogicalOperatorType::LOGICAL_PROJECTION) {
		projection = &child->Cast<LogicalProjection>();
		if (projection->children.size() != 1) {
			return nullptr;
		}
		child = projection->children[0].get();
	}
	if (child->type != LogicalOperatorType::LOGICAL_GET) {
		return nullptr;
	}
	auto &get = child->Cast<LogicalGet>();
	if (get.function.name != "read_fls" || !get.bind_data) {
		return nullptr;
	}
	auto &multi = get.bind_data->Cast<MultiFileBindData>();
	if (!multi.file_list || multi.file_list->GetExpandResult() != FileExpandResult::SINGLE_FILE || !multi.initial_reader ||
	    multi.file_options.union_by_name) {
		return nullptr;
	}
	auto &reader = multi.initial_reader->Cast<FastLanesReader>();
	const auto &descs = reader.GetColumnDescs();
	const auto get_bindings = get.GetColumnBindings();
	const auto &column_ids = get.GetColumnIds();
	// output position of the GET -> physical column of the file (by name and type: the file's own schema), or invalid
	auto physical_of_primary = [&](idx_t primary) -> optional_idx {
		if (primary >= get.names.size() || primary >= descs.size() || primary >= reader.columns.size()) {
			return optional_idx();
		}
		if (reader.columns[primary].name != get.names[primary] || reader.columns[primary].type != get.returned_types[primary]) {
			return optional_idx();
		}
		return primary;
	};
	auto physical_of_binding = [&](const ColumnBinding &binding) -> optional_idx {
		for (idx_t i = 0; i < get_bindings.size(); i++) {
			if (get_bindings[i] == binding) {
				const idx_t k = get.projection_ids.empty() ? i : get.projection_ids[i];
				if (k >= column_ids.size() || column_ids[k].IsVirtualColumn() || column_ids[k].HasChildren()) {
					return optional_idx();
				}
				return physical_of_primary(column_ids[k].GetPrimaryIndex());
			}
		}
		return optional_idx();
	};
	auto physical_of_expression = [&](Expression &expr) -> optional_idx {
		if (expr.GetExpressionClass() != ExpressionClass::BOUND_COLUMN_REF) {
			return optional_idx();
		}
		auto binding = expr.Cast<BoundColumnRefExpression>().binding;
		if (projection) {
			if (binding.table_index != projection->table_index || binding.column_index >= projection->expressions.size()) {
				return optional_idx();
			}
			auto &inner = *projection->expressions[binding.column_index];
			if (inner.GetExpressionClass() != ExpressionClass::BOUND_COLUMN_REF) {
				return optional_idx();
			}
			binding = inner.Cast<BoundColumnRefExpression>().binding;
		}
		return physical_of_binding(binding);
	};

	auto bind = make_uniq<FlsAggregateBindData>();
	// filters: everything the scan was asked to evaluate must go to the device, or the plan stays as it is
	for (auto &entry : get.table_filters.filters) {
		const auto physical = physical_of_primary(entry.first);
		if (!physical.IsValid()) {
			return nullptr;
		}
		const auto col = physical.GetIndex();
		if (!FastLanesFilterPushdown::ToDeviceFilter(*entry.second, descs[col], reader.columns[col].type, NumericCast<uint32_t>(col),
		                                             bind->filters, bind->filter_strings)) {
			return nullptr;
		}
	}
	bool has_extremes = false;
	vector<LogicalType> types;
	vector<string> names;
	for (auto &expr : aggr.expressions) {
		if (expr->GetExpressionClass() != ExpressionClass::BOUND_AGGREGATE) {
			return nullptr;
		}
		auto &aggregate = expr->Cast<BoundAggregateExpression>();
		if (aggregate.IsDistinct() || aggregate.filter || aggregate.order_bys) {
			return nullptr;
		}
		FlsAggregateSpec spec;
		spec.return_type = aggregate.return_type;
		const auto &fname = aggregate.function.name;
		if (fname == "count_star" && aggregate.children.empty()) {
			spec.kind = FlsAggregateKind::COUNT;
		} else if (aggregate.children.size() == 1) {
			const auto physical = physical_of_expression(*aggregate.children[0]);
			if (!physical.IsValid()) {
				return nullptr;
			}
			spec.column = NumericCast<uint32_t>(physical.GetIndex());
			const auto &column_type = reader.columns[spec.column].type;
			if (fname == "count") {
				spec.kind = FlsAggregateKind::COUNT; // no NULLs in a FastLanes file (fls_reader.cpp:199,208)
			} else if ((fname == "sum" || fname == "sum_no_overflow") && SummableType(column_type) &&
			           (spec.return_type.id() == LogicalTypeId::HUGEINT ||
			            (spec.return_type.id() == LogicalTypeId::DECIMAL && spec.return_type.InternalType() == PhysicalType::INT128))) {
				spec.kind = FlsAggregateKind::SUM;
			} else if ((fname == "min" || fname == "max") && OrderedType(column_type) && spec.return_type == column_type) {
				spec.kind = fname == "min" ? FlsAggregateKind::MIN : FlsAggregateKind::MAX;
				has_extremes = true;
			} else {
				return nullptr;
			}
		} else {
			return nullptr;
		}
		if (spec.kind == FlsAggregateKind::COUNT && spec.return_type.id() != LogicalTypeId::BIGINT) {
			return nullptr;
		}
		types.push_back(spec.return_type);
		names.push_back(fname + "_" + to_string(names.size()));
		bind->aggregates.push_back(std::move(spec));
	}
	// COUNT(*) WHERE ... without an aggregated column: the narrowest numeric column of the file carries the count
	for (idx_t c = 0; c < descs.size(); c++) {
		if (descs[c].phys_bits != 128 && (!bind->count_carrier.IsValid() || descs[c].phys_bits < descs[bind->count_carrier.GetIndex()].phys_bits)) {
			bind->count_carrier = c;
		}
	}
	if (!bind->count_carrier.IsValid() && !bind->filters.empty()) {
		return nullptr;
	}
	bind->table = reader.CachedTable();
	bind->path = reader.GetFileName();
	bind->n_rowgroups = NumericCast<uint32_t>(reader.GetNRowGroups());
	bind->n_tuples = reader.GetTotalTuples();
	if (has_extremes) {
		// zone maps answer MIN / MAX: only without filters and when every chunk of the column carries them
		if (!bind->filters.empty()) {
			return nullptr;
		}
		for (auto &spec : bind->aggregates) {
			if (spec.kind != FlsAggregateKind::MIN && spec.kind != FlsAggregateKind::MAX) {
				continue;
			}
			for (uint32_t r = 0; r < bind->n_rowgroups; r++) {
				fls_chunk_desc chunk;
				FlsCheck(flsgpu_table_chunk(reader.Table(), r, spec.column, &chunk), "flsgpu_table_chunk");
				if (!chunk.has_stats) {
					return nullptr;
				}
			}
		}
	}
	// string constants: the entries point at the bind data's copies
	for (auto &f : bind->filters) {
		if (descs[f.col].phys_bits == 128) {
			auto &bytes = bind->filter_strings[f.constant];
			f.str_data = reinterpret_cast<const uint8_t *>(bytes.data());
			f.str_len = bytes.size();
			f.constant = 0;
		}
	}
	TableFunction function("fls_aggregate_scan", {}, FlsAggregateScan, nullptr, FlsAggregateInit);
	function.to_string = FlsAggregateToString;
	const auto n_columns = types.size();
	auto result = make_uniq<LogicalGet>(aggr.aggregate_index, std::move(function), std::move(bind), std::move(types), std::move(names));
	for (idx_t i = 0; i < n_columns; i++) {
		result->AddColumnId(i);
	}
	result->SetEstimatedCardinality(1);
	(void)context;
	return std::move(result);
}

void RewriteAggregates(ClientContext &context, unique_ptr<LogicalOperator> &op) {
	for (auto &child : op->children) {
		RewriteAggregates(context, child);
	}
	if (op->type != LogicalOperatorType::LOGICAL_AGGREGATE_AND_GROUP_BY) {
		return;
	}
	auto replacement = TryRewriteAggregate(context, op->Cast<LogicalAggregate>());
	if (replacement) {
		op = std::move(replacement);
	}
}

void FlsOptimize(OptimizerExtensionInput &input, unique_ptr<LogicalOperator> &plan) {
	Value setting;
	if (input.context.TryGetCurrentSetting("fastlanes_aggregate_pushdown", setting) && !setting.IsNull() && !setting.GetValue<bool>()) {
		return;
	}
	RewriteAggregates(input.context, plan);
}

} // namespace

void FastLanesAggregatePushdown::Register(ExtensionLoader &loader) {
	auto &config = DBConfig::GetConfig(loader.GetDatabaseInstance());
	config.AddExtensionOption("fastlanes_aggregate_pushdown",
	                          "Answer ungrouped SUM / COUNT / MIN / MAX over read_fls columns (with constant-comparison filters) by "
	                          "fused decode + aggregate kernels instead of a materialising scan",
	                          LogicalType::BOOLEAN, Value::BOOLEAN(true));
	OptimizerExtension extension;
	extension.optimize_function = FlsOptimize;
	config.optimizer_extensions.push_back(std::move(extension));
}

} // namespace duckdb
