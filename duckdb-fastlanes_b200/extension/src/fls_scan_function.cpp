// fls_scan_function.cpp — read_fls: registration and the MultiFileReaderInterface glue.
// Same surface as the reference (src/read_fls.cpp:9-46, src/reader/fls_multi_file_info.cpp:7-171): named parameters
// explicit_cardinality / file_row_number, statistics, row-id columns, partition stats, filter_pushdown + filter_prune,
// no late materialization, VARCHAR and LIST(VARCHAR) overloads through MultiFileReader::CreateFunctionSet.
#include "fastlanes_extension.hpp"
#include "fls_file_reader.hpp"

#include "duckdb/common/multi_file/multi_file_reader.hpp"
#include "duckdb/common/string_util.hpp"
#include "duckdb/main/extension/extension_loader.hpp"
#include "duckdb/parallel/task_scheduler.hpp"

namespace duckdb {

// ---- FastLanesMultiFileInfo ---------------------------------------------------------------------------------------

unique_ptr<MultiFileReaderInterface> FastLanesMultiFileInfo::CreateInterface(ClientContext &) {
	return make_uniq<FastLanesMultiFileInfo>();
}

unique_ptr<BaseFileReaderOptions> FastLanesMultiFileInfo::InitializeOptions(ClientContext &, optional_ptr<TableFunctionInfo>) {
	return make_uniq<FastLanesFileReaderOptions>();
}

bool FastLanesMultiFileInfo::ParseCopyOption(ClientContext &, const string &, const vector<Value> &, BaseFileReaderOptions &,
                                             vector<string> &, vector<LogicalType> &) {
	return false; // COPY ... FROM has no fls-specific options (fls_multi_file_info.cpp:11-18)
}

bool FastLanesMultiFileInfo::ParseOption(ClientContext &, const string &original_key, const Value &val, MultiFileOptions &,
                                         BaseFileReaderOptions &options_p) {
	auto &options = options_p.Cast<FastLanesFileReaderOptions>();
	const auto key = StringUtil::Lower(original_key);
	if (val.IsNull()) {
		throw BinderException("Cannot use NULL as argument to %s", original_key);
	}
	if (key == "explicit_cardinality") {
		options.explicit_cardinality = UBigIntValue::Get(val);
		return true;
	}
	if (key == "file_row_number") {
		options.file_row_number = BooleanValue::Get(val);
		return true;
	}
	return false;
}

unique_ptr<TableFunctionData> FastLanesMultiFileInfo::InitializeBindData(MultiFileBindData &multi_file_data,
                                                                         unique_ptr<BaseFileReaderOptions> options_p) {
	auto bind_data = make_uniq<FastLanesReadBindData>();
	bind_data->options = unique_ptr_cast<BaseFileReaderOptions, FastLanesFileReaderOptions>(std::move(options_p));
	if (bind_data->options->explicit_cardinality) {
		const auto file_count = multi_file_data.file_list->GetTotalFileCount();
		bind_data->initial_file_cardinality = bind_data->options->explicit_cardinality / (file_count ? file_count : 1);
	}
	return std::move(bind_data);
}

void FastLanesMultiFileInfo::BindReader(ClientContext &context, vector<LogicalType> &return_types, vector<string> &names,
                                        MultiFileBindData &bind_data) {
	auto &bind = bind_data.bind_data->Cast<FastLanesReadBindData>();
	if (bind_data.file_options.union_by_name) {
		bind_data.reader_bind = bind_data.multi_file_reader->BindUnionReader(context, return_types, names, *bind_data.file_list,
		                                                                     bind_data, *bind.options, bind_data.file_options);
	} else {
		bind_data.reader_bind = bind_data.multi_file_reader->BindReader(context, return_types, names, *bind_data.file_list, bind_data,
		                                                                *bind.options, bind_data.file_options);
	}
}

void FastLanesMultiFileInfo::FinalizeBindData(MultiFileBindData &multi_file_data) {
	auto &bind_data = multi_file_data.bind_data->Cast<FastLanesReadBindData>();
	bind_data.initial_file_cardinality = 0;
	bind_data.initial_file_n_rowgroups = 0;
	if (multi_file_data.initial_reader) {
		auto &initial_reader = multi_file_data.initial_reader->Cast<FastLanesReader>();
		bind_data.initial_file_cardinality = initial_reader.GetTotalTuples();
		bind_data.initial_file_n_rowgroups = initial_reader.GetNRowGroups();
		if (bind_data.options) {
			bind_data.options->file_row_number = initial_reader.GetOptions().file_row_number;
		}
	}
}

void FastLanesMultiFileInfo::GetBindInfo(const TableFunctionData &bind_data_p, BindInfo &info) {
	// the reference casts the bind data to its options class here (fls_multi_file_info.cpp:48-55); the
	// options live INSIDE the bind data
	auto &bind_data = bind_data_p.Cast<FastLanesReadBindData>();
	if (bind_data.options) {
		info.InsertOption("file_row_number", Value::BOOLEAN(bind_data.options->file_row_number));
		info.InsertOption("explicit_cardinality", Value::UBIGINT(bind_data.options->explicit_cardinality));
	}
}

optional_idx FastLanesMultiFileInfo::MaxThreads(const MultiFileBindData &bind_data_p, const MultiFileGlobalState &,
                                                FileExpandResult expand_result) {
	// several files: as many threads as DuckDB has, so that a small first file does not serialise the scan
	// (fls_multi_file_info.cpp:99-110); one file: one worker per row group at most
	if (expand_result == FileExpandResult::MULTIPLE_FILES) {
		return optional_idx();
	}
	const auto &bind_data = bind_data_p.bind_data->Cast<FastLanesReadBindData>();
	return MaxValue<idx_t>(bind_data.initial_file_n_rowgroups, 1);
}

unique_ptr<GlobalTableFunctionState> FastLanesMultiFileInfo::InitializeGlobalState(ClientContext &, MultiFileBindData &,
                                                                                   MultiFileGlobalState &) {
	return make_uniq<FastLanesReadGlobalState>();
}

unique_ptr<LocalTableFunctionState> FastLanesMultiFileInfo::InitializeLocalState(ExecutionContext &, GlobalTableFunctionState &) {
	return make_uniq<FastLanesReadLocalState>();
}

shared_ptr<BaseFileReader> FastLanesMultiFileInfo::CreateReader(ClientContext &context, GlobalTableFunctionState &,
                                                                BaseUnionData &union_data_p, const MultiFileBindData &) {
	auto &union_data = union_data_p.Cast<FastLanesUnionData>();
	return make_shared_ptr<FastLanesReader>(context, union_data.file, union_data.options);
}

shared_ptr<BaseFileReader> FastLanesMultiFileInfo::CreateReader(ClientContext &context, GlobalTableFunctionState &,
                                                                const OpenFileInfo &file, idx_t, const MultiFileBindData &bind_data_p) {
	const auto &bind_data = bind_data_p.bind_data->Cast<FastLanesReadBindData>();
	return make_shared_ptr<FastLanesReader>(context, file, *bind_data.options);
}

shared_ptr<BaseFileReader> FastLanesMultiFileInfo::CreateReader(ClientContext &context, const OpenFileInfo &file,
                                                                BaseFileReaderOptions &options_p, const MultiFileOptions &) {
	return make_shared_ptr<FastLanesReader>(context, file, options_p.Cast<FastLanesFileReaderOptions>());
}

void FastLanesMultiFileInfo::FinishReading(ClientContext &, GlobalTableFunctionState &, LocalTableFunctionState &) {
}

unique_ptr<NodeStatistics> FastLanesMultiFileInfo::GetCardinality(const MultiFileBindData &bind_data_p, idx_t file_count) {
	const auto &bind_data = bind_data_p.bind_data->Cast<FastLanesReadBindData>();
	if (const auto explicit_cardinality = bind_data.options->explicit_cardinality) {
		return make_uniq<NodeStatistics>(explicit_cardinality, explicit_cardinality);
	}
	if (file_count == 1 && bind_data.initial_file_cardinality) {
		return make_uniq<NodeStatistics>(bind_data.initial_file_cardinality, bind_data.initial_file_cardinality);
	}
	// estimate from the first file; 42 rows per file when it is empty (fls_multi_file_info.cpp:152-164)
	return make_uniq<NodeStatistics>(MaxValue<idx_t>(bind_data.initial_file_cardinality, 42) * file_count);
}

void FastLanesMultiFileInfo::GetVirtualColumns(ClientContext &, MultiFileBindData &, virtual_column_map_t &result) {
	result.insert(
	    make_pair(MultiFileReader::COLUMN_IDENTIFIER_FILE_ROW_NUMBER, TableColumn("file_row_number", LogicalType::BIGINT)));
}

// ---- table function -----------------------------------------------------------------------------------------------

static bool FastLanesPushdownExpression(ClientContext &, const LogicalGet &, Expression &) {
	return true;
}

static vector<column_t> FastLanesGetRowIdColumns(ClientContext &, optional_ptr<FunctionData>) {
	return {MultiFileReader::COLUMN_IDENTIFIER_FILE_INDEX, MultiFileReader::COLUMN_IDENTIFIER_FILE_ROW_NUMBER};
}

// COUNT(*) and row-id range pruning are answered from the footer (read_fls.cpp:20-30, fls_reader.cpp:316-332)
static vector<PartitionStatistics> FastLanesGetPartitionStats(ClientContext &, GetPartitionStatsInput &input) {
	auto &bind_data = input.bind_data->Cast<MultiFileBindData>();
	vector<PartitionStatistics> result;
	if (bind_data.file_list->GetExpandResult() == FileExpandResult::SINGLE_FILE && bind_data.initial_reader) {
		bind_data.initial_reader->Cast<FastLanesReader>().GetPartitionStats(result);
	}
	return result;
}

void FastLanesScanFunction::Register(ExtensionLoader &loader) {
	MultiFileFunction<FastLanesMultiFileInfo> fn("read_fls");
	fn.named_parameters["explicit_cardinality"] = LogicalType::UBIGINT;
	fn.named_parameters["file_row_number"] = LogicalType::BOOLEAN;
	fn.statistics = MultiFileFunction<FastLanesMultiFileInfo>::MultiFileScanStats;
	fn.get_row_id_columns = FastLanesGetRowIdColumns;
	fn.pushdown_expression = FastLanesPushdownExpression;
	fn.get_partition_stats = FastLanesGetPartitionStats;
	fn.filter_pushdown = true;
	fn.filter_prune = true;
	fn.late_materialization = false;
	loader.RegisterFunction(MultiFileReader::CreateFunctionSet(static_cast<TableFunction>(fn)));
}

} // namespace duckdb
