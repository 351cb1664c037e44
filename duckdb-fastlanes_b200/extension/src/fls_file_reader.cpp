// fls_file_reader.cpp — FastLanesReader : BaseFileReader over libflsgpu (see fls_file_reader.hpp).
// Each method names the reference method it stands in for (reference src/reader/fls_reader.cpp).
#include "fls_file_reader.hpp"

#include "duckdb/common/multi_file/multi_file_reader.hpp"
#include "duckdb/common/string_util.hpp"
#include "duckdb/common/types/selection_vector.hpp"
#include "duckdb/execution/expression_executor.hpp"
#include "duckdb/parallel/task_scheduler.hpp"
#include "duckdb/storage/statistics/numeric_stats.hpp"
#include "duckdb/storage/statistics/string_stats.hpp"
#include "duckdb/storage/table/column_segment.hpp"
#include "duckdb/storage/table/row_group.hpp"

#include <cstring>
#include <sys/stat.h>

namespace duckdb {

void FlsCheck(int status, const char *where) {
	if (status == FLSGPU_OK) {
		return;
	}
	const string msg = string(where) + ": " + flsgpu_last_error();
	switch (status) {
	case FLSGPU_ERR_IO:
		throw IOException(msg);
	case FLSGPU_ERR_UNSUPPORTED:
		throw NotImplementedException(msg);
	case FLSGPU_ERR_CONVERSION:
		throw ConversionException(msg);
	case FLSGPU_ERR_CUDA:
		// no CPU fallback: without a usable device the scan fails, loudly
		throw IOException("FastLanes GPU engine: " + msg);
	default:
		throw InvalidInputException(msg);
	}
}

FlsPinnedBuffer::FlsPinnedBuffer(idx_t bytes_p) : VectorBuffer(VectorBufferType::OPAQUE_BUFFER), bytes(bytes_p) {
	void *p = nullptr;
	FlsCheck(flsgpu_host_alloc(bytes_p, &p), "flsgpu_host_alloc");
	ptr = data_ptr_cast(p);
}

FlsPinnedBuffer::~FlsPinnedBuffer() {
	flsgpu_host_free(ptr);
}

// A staging buffer of at least `bytes`: the current one when nobody downstream still points into it, else a new one.
static void ReservePinned(buffer_ptr<FlsPinnedBuffer> &buf, idx_t bytes) {
	if (buf && buf->bytes >= bytes && buf.use_count() == 1) {
		return;
	}
	buf = make_buffer<FlsPinnedBuffer>(MaxValue<idx_t>(bytes, 16));
}

// TranslateUtils::TranslateType (reference src/reader/translation_utils.cpp:5-48) + the DECIMAL annotation
// (fls_reader.cpp:119-126).  out_type = the DuckDB physical type the decode kernel writes: every unsigned width is
// widened to UBIGINT, everything else keeps its width.
static LogicalType TranslateColumnType(const fls_column_desc &col, int32_t &out_type) {
	out_type = FLSGPU_OUT_NATIVE;
	switch (col.data_type) {
	case FLS_TYPE_INT8:
		return LogicalType::TINYINT;
	case FLS_TYPE_INT16:
		return LogicalType::SMALLINT;
	case FLS_TYPE_INT32:
		return LogicalType::INTEGER;
	case FLS_TYPE_INT64:
		return LogicalType::BIGINT;
	case FLS_TYPE_UINT8:
	case FLS_TYPE_UINT16:
	case FLS_TYPE_UINT32:
	case FLS_TYPE_UINT64:
		out_type = FLSGPU_OUT_U64;
		return LogicalType::UBIGINT;
	case FLS_TYPE_FLOAT:
		return LogicalType::FLOAT;
	case FLS_TYPE_DOUBLE:
		return LogicalType::DOUBLE;
	case FLS_TYPE_BOOLEAN:
		return LogicalType::BOOLEAN;
	case FLS_TYPE_DATE:
		return LogicalType::DATE;
	case FLS_TYPE_TIMESTAMP:
		return LogicalType::TIMESTAMP;
	case FLS_TYPE_DECIMAL:
		return LogicalType::DECIMAL(col.decimal_precision, col.decimal_scale);
	case FLS_TYPE_FLS_STR:
		return LogicalType::VARCHAR;
	case FLS_TYPE_BYTE_ARRAY:
		return LogicalType::BLOB;
	default:
		throw InternalException("TranslateType: column type is not supported");
	}
}

FastLanesReader::FastLanesReader(ClientContext &context, OpenFileInfo file_p, const FastLanesFileReaderOptions &options)
    : BaseFileReader(std::move(file_p)), vectors_read(0), reader_options(options), rowgroup_filter_ready(false) {
	Initialize(context);
	if (reader_options.file_row_number) {
		EnsureFileRowNumberColumn();
	}
}

FastLanesReader::~FastLanesReader() {
	// the handles belong to cached_table: closed when the last reader and the cache let go of the entry
}

// SET fastlanes_gpu_devices = '0,1,2,3': the process-wide device list of the engine (flsgpu_init); applied when it changes
static void ApplyDeviceList(const string &list) {
	static std::mutex lock;
	static string applied;
	std::lock_guard<std::mutex> guard(lock);
	if (list == applied) {
		return;
	}
	vector<int> ordinals;
	for (auto &part : StringUtil::Split(list, ',')) {
		auto text = part;
		StringUtil::Trim(text);
		if (text.empty()) {
			continue;
		}
		for (auto ch : text) {
			if (ch < '0' || ch > '9') {
				throw InvalidInputException("fastlanes_gpu_devices: \"%s\" is not a list of CUDA device ordinals", list);
			}
		}
		ordinals.push_back(std::stoi(text));
	}
	FlsCheck(flsgpu_init(NumericCast<int>(ordinals.size()), ordinals.data()), "fastlanes_gpu_devices");
	applied = list;
}

// ---- FlsTableCache ------------------------------------------------------------------------------------------------------
FlsCachedTable::~FlsCachedTable() {
	for (idx_t i = handles.size(); i-- > 1;) {
		flsgpu_table_close(handles[i]);
	}
	if (!handles.empty()) {
		flsgpu_table_close(handles[0]);
	}
}

shared_ptr<FlsCachedTable> FlsTableCache::Open(const string &path, const string &key) {
	auto entry = make_shared_ptr<FlsCachedTable>();
	entry->key = key;
	flsgpu_table *table = nullptr;
	FlsCheck(flsgpu_table_open(path.c_str(), &table), "read_fls");
	entry->handles.push_back(table);
	flsgpu_table_info info {};
	FlsCheck(flsgpu_table_info_get(table, &info), "flsgpu_table_info_get");
	entry->file_bytes = info.file_bytes;
	int n_devices = 1;
	FlsCheck(flsgpu_n_devices(&n_devices), "flsgpu_n_devices");
	for (int g = 1; g < n_devices; g++) {
		flsgpu_table *shared = nullptr;
		FlsCheck(flsgpu_table_share(table, &shared), "flsgpu_table_share");
		entry->handles.push_back(shared);
		FlsCheck(flsgpu_table_bind_device(shared, g), "flsgpu_table_bind_device");
	}
	return entry;
}

shared_ptr<FlsCachedTable> FlsTableCache::Acquire(ClientContext &context, const string &path) {
	static std::mutex lock;
	// most recently used first; never destroyed: at process exit the CUDA runtime may be gone before static destructors run
	static auto &entries = *new vector<shared_ptr<FlsCachedTable>>();
	string device_list;
	Value setting;
	if (context.TryGetCurrentSetting("fastlanes_gpu_devices", setting) && !setting.IsNull()) {
		device_list = setting.ToString();
	}
	ApplyDeviceList(device_list); // fails here, at bind time, when there is no such device
	idx_t capacity_mb = 65536;
	if (context.TryGetCurrentSetting("fastlanes_table_cache_mb", setting) && !setting.IsNull()) {
		capacity_mb = setting.GetValue<uint64_t>();
	}
	struct stat st {};
	if (capacity_mb == 0 || stat(path.c_str(), &st) != 0) {
		return Open(path, string()); // not cached (or not a local file: flsgpu_table_open reports what is wrong with it)
	}
	const string key = path + "|" + to_string((uint64_t)st.st_size) + "|" + to_string((uint64_t)st.st_mtim.tv_sec) + "." +
	                   to_string((uint64_t)st.st_mtim.tv_nsec) + "|" + device_list;
	auto lookup = [&]() -> shared_ptr<FlsCachedTable> {
		for (idx_t i = 0; i < entries.size(); i++) {
			if (entries[i]->key == key) {
				auto hit = entries[i];
				entries.erase(entries.begin() + NumericCast<int64_t>(i));
				entries.insert(entries.begin(), hit);
				return hit;
			}
		}
		return nullptr;
	};
	{
		std::lock_guard<std::mutex> guard(lock);
		if (auto hit = lookup()) {
			return hit;
		}
	}
	// reading and validating a file takes its time (0.08 s per 200 MB): not under the lock, the files of a glob open in parallel
	auto entry = Open(path, key);
	std::lock_guard<std::mutex> guard(lock);
	if (auto hit = lookup()) {
		return hit; // another thread opened the same file meanwhile: one copy is enough
	}
	// an older version of the same file is of no use to anybody new
	const string prefix = path + "|";
	for (idx_t i = entries.size(); i-- > 0;) {
		if (StringUtil::StartsWith(entries[i]->key, prefix)) {
			entries.erase(entries.begin() + NumericCast<int64_t>(i));
		}
	}
	entries.insert(entries.begin(), entry);
	idx_t total = 0;
	for (auto &e : entries) {
		total += e->file_bytes;
	}
	for (idx_t i = entries.size(); i-- > 1 && total > capacity_mb * 1024 * 1024;) {
		if (entries[i].use_count() == 1) { // nobody scans it
			total -= entries[i]->file_bytes;
			entries.erase(entries.begin() + NumericCast<int64_t>(i));
		}
	}
	return entry;
}

// FastLanesReader::Initialize (fls_reader.cpp:106-143): footer -> schema, per-row-group tuple offsets, totals
void FastLanesReader::Initialize(ClientContext &context) {
	cached_table = FlsTableCache::Acquire(context, file.path);
	device_tables = cached_table->handles;
	table = device_tables[0];
	FlsCheck(flsgpu_table_info_get(table, &info), "flsgpu_table_info_get");
	columns.clear();
	column_descs.resize(info.n_columns);
	column_out_types.resize(info.n_columns);
	for (uint32_t c = 0; c < info.n_columns; c++) {
		FlsCheck(flsgpu_table_column(table, c, &column_descs[c]), "flsgpu_table_column");
		const auto &desc = column_descs[c];
		columns.emplace_back(string(desc.name, strnlen(desc.name, sizeof(desc.name))), TranslateColumnType(desc, column_out_types[c]));
	}
	row_groups.resize(info.n_rowgroups);
	row_group_offsets.resize(info.n_rowgroups);
	for (uint32_t r = 0; r < info.n_rowgroups; r++) {
		FlsCheck(flsgpu_table_rowgroup(table, r, &row_groups[r]), "flsgpu_table_rowgroup");
		row_group_offsets[r] = total_tuples;
		total_tuples += row_groups[r].n_tuples;
		total_vectors += row_groups[r].n_vec;
	}
	Value setting;
	if (context.TryGetCurrentSetting("fastlanes_rowgroups_per_batch", setting) && !setting.IsNull()) {
		rowgroups_per_batch = MaxValue<idx_t>(1, setting.GetValue<uint64_t>());
	}
	if (context.TryGetCurrentSetting("fastlanes_compressed_vectors", setting) && !setting.IsNull()) {
		compressed_vectors = setting.GetValue<bool>();
	}
	if (context.TryGetCurrentSetting("fastlanes_select_on_device", setting) && !setting.IsNull()) {
		select_on_device = setting.GetValue<bool>();
	}
}

void FastLanesReader::EnsureFileRowNumberColumn() {
	if (file_row_number_local_idx.IsValid()) {
		return;
	}
	MultiFileColumnDefinition result("file_row_number", LogicalType::BIGINT);
	result.identifier = Value::INTEGER(MultiFileReader::ORDINAL_FIELD_ID);
	file_row_number_local_idx = columns.size();
	columns.push_back(std::move(result));
	reader_options.file_row_number = true;
}

bool FastLanesReader::IsFileRowNumberColumn(column_t column_id) const {
	return file_row_number_local_idx.IsValid() && column_id == file_row_number_local_idx.GetIndex();
}

void FastLanesReader::AddVirtualColumn(column_t virtual_column_id) {
	if (virtual_column_id == MultiFileReader::COLUMN_IDENTIFIER_FILE_ROW_NUMBER) {
		EnsureFileRowNumberColumn();
		return;
	}
	throw InternalException("Unsupported virtual column id %d for FastLanes reader", virtual_column_id);
}

shared_ptr<BaseUnionData> FastLanesReader::GetUnionData(idx_t file_idx) {
	auto data = make_shared_ptr<FastLanesUnionData>(file);
	for (auto &column : columns) {
		data->names.push_back(column.name);
		data->types.push_back(column.type);
	}
	data->options = reader_options;
	if (file_idx == 0) {
		data->reader = shared_from_this();
	}
	return std::move(data);
}

// ---- statistics (fls_reader.cpp:187-292, row_group_statistics.cpp:60-127) -----------------------------------------

// zone map word of a chunk -> Value of the column's logical type
static bool ZoneMapValue(const fls_column_desc &desc, const LogicalType &type, uint64_t word, Value &out) {
	if (desc.is_float) {
		double d;
		std::memcpy(&d, &word, sizeof(d));
		if (Value::IsNan(d)) {
			return false;
		}
		out = type.id() == LogicalTypeId::FLOAT ? Value::FLOAT(static_cast<float>(d)) : Value::DOUBLE(d);
		return true;
	}
	if (type.id() == LogicalTypeId::UBIGINT) {
		out = Value::UBIGINT(word);
		return true;
	}
	if (type.id() == LogicalTypeId::BOOLEAN) {
		out = Value::BOOLEAN(word != 0);
		return true;
	}
	out = Value::Numeric(type, static_cast<int64_t>(word));
	return true;
}

unique_ptr<BaseStatistics> FastLanesReader::GetStatistics(ClientContext &, const string &name) {
	idx_t file_col_idx;
	for (file_col_idx = 0; file_col_idx < columns.size(); file_col_idx++) {
		if (columns[file_col_idx].name == name) {
			break;
		}
	}
	if (file_col_idx == columns.size()) {
		return nullptr;
	}
	const auto &type = columns[file_col_idx].type;
	if (IsFileRowNumberColumn(file_col_idx)) {
		auto stats = NumericStats::CreateUnknown(type);
		NumericStats::SetMin(stats, Value::BIGINT(0));
		NumericStats::SetMax(stats, Value::BIGINT(NumericCast<int64_t>(total_tuples)));
		stats.Set(StatsInfo::CANNOT_HAVE_NULL_VALUES); // FastLanes files hold no NULLs
		return stats.ToUnique();
	}
	if (type.id() == LogicalTypeId::VARCHAR) {
		// dictionary-encoded columns: longest string, min / max prefix and the ASCII flag come from the dictionaries (a few
		// strings per row group) - what DuckDB's compressed materialization needs to turn GROUP BY l_returnflag, l_linestatus
		// into a perfect hash aggregate over one-byte keys, as it does over its own tables
		flsgpu_string_stats summary {};
		if (file_col_idx < info.n_columns && flsgpu_column_string_stats(table, NumericCast<uint32_t>(file_col_idx), &summary) == FLSGPU_OK &&
		    summary.available) {
			auto stats = StringStats::CreateEmpty(type);
			StringStats::Update(stats, string_t(const_char_ptr_cast(summary.min_value), summary.min_length));
			StringStats::Update(stats, string_t(const_char_ptr_cast(summary.max_value), summary.max_value_length));
			StringStats::SetMaxStringLength(stats, summary.max_length);
			if (summary.has_non_ascii) {
				StringStats::SetContainsUnicode(stats);
			}
			stats.Set(StatsInfo::CANNOT_HAVE_NULL_VALUES);
			return stats.ToUnique();
		}
		auto stats = StringStats::CreateUnknown(type);
		stats.Set(StatsInfo::CANNOT_HAVE_NULL_VALUES);
		return stats.ToUnique();
	}
	if (!type.IsNumeric() && type.id() != LogicalTypeId::DATE && type.id() != LogicalTypeId::TIMESTAMP &&
	    type.id() != LogicalTypeId::BOOLEAN) {
		return nullptr;
	}
	const auto &desc = column_descs[file_col_idx];
	unique_ptr<BaseStatistics> column_stats;
	for (uint32_t r = 0; r < info.n_rowgroups; r++) {
		fls_chunk_desc chunk;
		FlsCheck(flsgpu_table_chunk(table, r, NumericCast<uint32_t>(file_col_idx), &chunk), "flsgpu_table_chunk");
		auto chunk_stats = NumericStats::CreateUnknown(type);
		Value min_val, max_val;
		if (chunk.has_stats && ZoneMapValue(desc, type, chunk.min_val, min_val) && ZoneMapValue(desc, type, chunk.max_val, max_val)) {
			NumericStats::SetMin(chunk_stats, min_val);
			NumericStats::SetMax(chunk_stats, max_val);
		}
		chunk_stats.Set(StatsInfo::CANNOT_HAVE_NULL_VALUES);
		if (!column_stats) {
			column_stats = chunk_stats.ToUnique();
		} else {
			column_stats->Merge(chunk_stats);
		}
	}
	return column_stats;
}

// FastLanesReader::GetPartitionStats (fls_reader.cpp:316-332): exact row ranges per row group
void FastLanesReader::GetPartitionStats(vector<PartitionStatistics> &result) const {
	result.reserve(info.n_rowgroups);
	for (uint32_t r = 0; r < info.n_rowgroups; r++) {
		PartitionStatistics partition_stats;
		partition_stats.row_start = row_group_offsets[r];
		partition_stats.count = row_groups[r].n_tuples;
		partition_stats.count_type = CountType::COUNT_EXACT;
		result.push_back(partition_stats);
	}
}

// ---- row-group hand-out ----------------------------------------------------------------------------------------------

// FastLanesReader::BuildReaderProjectionPlan (fls_reader.cpp:560-613).  The multi-column-compression fallback (decode every
// column of the row group when a projected one is EXP_EQUAL, :43-82,582-590) is not needed: an EXP_EQUAL chunk of this
// container carries its source chunk's descriptor, so it stages and decodes like any other chunk.
FastLanesReaderProjectionPlan FastLanesReader::BuildReaderProjectionPlan() {
	FastLanesReaderProjectionPlan plan;
	plan.output_to_reader_idx.resize(column_ids.size());
	for (idx_t i = 0; i < column_ids.size(); i++) {
		const auto col_idx = column_ids[MultiFileLocalIndex(i)].GetId();
		if (IsFileRowNumberColumn(col_idx)) {
			continue;
		}
		if (col_idx >= info.n_columns) {
			throw InternalException("Projected column index out of range");
		}
		idx_t reader_idx;
		for (reader_idx = 0; reader_idx < plan.reader_column_ids.size(); reader_idx++) {
			if (plan.reader_column_ids[reader_idx] == col_idx) {
				break;
			}
		}
		if (reader_idx == plan.reader_column_ids.size()) {
			plan.reader_column_ids.push_back(NumericCast<uint32_t>(col_idx));
		}
		plan.output_to_reader_idx[i] = reader_idx;
	}
	plan.has_physical_columns = !plan.reader_column_ids.empty();
	return plan;
}

// RowGroupFilter::EnsureRowGroups / RowGroupMaySatisfyFilters (row_group_filter.cpp:62-199): zone maps against the constant
// comparisons the device evaluates anyway (flsgpu_rowgroup_prune), row-id ranges against filters on file_row_number.
const vector<idx_t> &FastLanesReader::GetRowGroupsToScan(ClientContext &context, FastLanesReadLocalState &local_state) {
	if (rowgroup_filter_ready.load(std::memory_order_acquire) && rowgroup_filter_for == filters.get()) {
		return rowgroups_to_scan;
	}
	std::lock_guard<std::mutex> guard(rowgroup_filter_lock);
	if (rowgroup_filter_ready.load(std::memory_order_acquire) && rowgroup_filter_for == filters.get()) {
		return rowgroups_to_scan;
	}
	rowgroups_to_scan.clear();
	vector<uint8_t> keep(info.n_rowgroups, 1);
	if (filters && !filters->filters.empty() && info.n_rowgroups) {
		InitializeScanFilters(context, local_state);
		if (!local_state.device_filters.empty()) {
			FlsCheck(flsgpu_rowgroup_prune(table, local_state.device_filters.data(), NumericCast<uint32_t>(local_state.device_filters.size()),
			                               0, info.n_rowgroups, keep.data()),
			         "flsgpu_rowgroup_prune");
		}
		for (auto &entry : filters->filters) {
			if (entry.first >= column_ids.size() || !IsFileRowNumberColumn(column_ids[MultiFileLocalIndex(entry.first)].GetId())) {
				continue;
			}
			for (uint32_t r = 0; r < info.n_rowgroups; r++) {
				const idx_t begin = row_group_offsets[r], end = begin + row_groups[r].n_tuples;
				if (keep[r] && RowGroup::CheckRowIdFilter(*entry.second, begin, end) == FilterPropagateResult::FILTER_ALWAYS_FALSE) {
					keep[r] = 0;
				}
			}
		}
	}
	for (uint32_t r = 0; r < info.n_rowgroups; r++) {
		if (keep[r]) {
			rowgroups_to_scan.push_back(r);
		}
	}
	rowgroup_filter_for = filters.get();
	rowgroup_filter_ready.store(true, std::memory_order_release);
	return rowgroups_to_scan;
}

// FastLanesReader::TryAssignNextRowGroup (fls_reader.cpp:337-355), handing out a run of consecutive row groups
bool FastLanesReader::TryAssignNextRowGroup(FastLanesReadGlobalState &global_state, FastLanesReadLocalState &local_state,
                                            idx_t batch_limit) {
	const auto &to_scan = rowgroups_to_scan;
	if (global_state.next_rowgroup >= to_scan.size()) {
		return false;
	}
	idx_t n = 1;
	while (n < batch_limit && global_state.next_rowgroup + n < to_scan.size() &&
	       to_scan[global_state.next_rowgroup + n] == to_scan[global_state.next_rowgroup] + n) {
		n++;
	}
	local_state.batch_begin = to_scan[global_state.next_rowgroup];
	local_state.batch_end = local_state.batch_begin + n;
	global_state.next_rowgroup += n; // mark the row groups as consumed (:511)
	local_state.device_slot = global_state.next_batch++ % device_tables.size();
	local_state.cur_rowgroup = local_state.batch_begin;
	local_state.projection = BuildReaderProjectionPlan();
	local_state.row_group_base = row_group_offsets[local_state.cur_rowgroup];
	local_state.cur_vector = 0;
	local_state.n_tuples = GetNTuples(local_state.cur_rowgroup);
	local_state.batch_decoded = false;
	return true;
}

// FastLanesReader::InitializeScanFilters (fls_reader.cpp:357-380): filters are equal across a file, classify them once
void FastLanesReader::InitializeScanFilters(ClientContext &context, FastLanesReadLocalState &local_state) {
	if (local_state.filters_ready && local_state.filters_for == this) {
		return;
	}
	local_state.device_filters.clear();
	local_state.device_filter_strings.clear();
	local_state.host_filters.clear();
	local_state.conversion_executor.reset();
	local_state.conversion_of_output.assign(column_ids.size(), optional_idx());
	if (filters) {
		FastLanesFilterPushdown::Split(context, *this, column_descs, info.n_columns, local_state);
	}
	// conversion expressions the multi-file reader moved into this reader (multi_file_column_mapper.cpp:1088-1099): the
	// output column then carries the expression's type, evaluated over the decoded column
	if (!expression_map.empty()) {
		vector<LogicalType> input_types;
		for (idx_t i = 0; i < column_ids.size(); i++) {
			const auto local_id = column_ids[MultiFileLocalIndex(i)].GetId();
			auto entry = expression_map.find(local_id);
			if (entry == expression_map.end()) {
				continue;
			}
			if (!local_state.conversion_executor) {
				local_state.conversion_executor = make_uniq<ExpressionExecutor>(context);
			}
			local_state.conversion_of_output[i] = local_state.conversion_executor->expressions.size();
			local_state.conversion_executor->AddExpression(*entry->second);
			input_types.push_back(columns[local_id].type);
		}
	}
	local_state.filters_ready = true;
	local_state.filters_for = this;
}

// FastLanesReader::TryInitializeScan (fls_reader.cpp:497-514).  Runs under the multi-file global lock
// (multi_file_function.hpp:399-416): only bookkeeping happens here, the GPU call is made by the first Scan of the batch.
bool FastLanesReader::TryInitializeScan(ClientContext &context, GlobalTableFunctionState &gstate_p, LocalTableFunctionState &lstate_p) {
	auto &global_state = gstate_p.Cast<FastLanesReadGlobalState>();
	auto &local_state = lstate_p.Cast<FastLanesReadLocalState>();
	GetRowGroupsToScan(context, local_state);
	InitializeScanFilters(context, local_state);
	// batch size: the setting, but never so large that DuckDB's other workers are left without row groups
	const idx_t n_threads = MaxValue<idx_t>(1, NumericCast<idx_t>(TaskScheduler::GetScheduler(context).NumberOfThreads()));
	const idx_t fair_share = (rowgroups_to_scan.size() + 4 * n_threads - 1) / (4 * n_threads);
	const idx_t batch_limit = MaxValue<idx_t>(1, MinValue<idx_t>(rowgroups_per_batch, fair_share));
	if (!TryAssignNextRowGroup(global_state, local_state, batch_limit)) {
		local_state.batch_begin = local_state.batch_end = 0;
		return false;
	}
	return true;
}

void FastLanesReader::FinishFile(ClientContext &, GlobalTableFunctionState &gstate_p) {
	gstate_p.Cast<FastLanesReadGlobalState>().next_rowgroup = 0; // progress trackers are per file (fls_reader.cpp:549-554)
}

double FastLanesReader::GetProgressInFile(ClientContext &) {
	return total_vectors ? 100.0 * (static_cast<double>(vectors_read.load()) / static_cast<double>(total_vectors)) : 100.0;
}

// ---- decode ----------------------------------------------------------------------------------------------------------

// get_rowgroup_reader + InitializeColumnDecoders + every Decode<PASS> of the batch (fls_reader.cpp:382-420,436-472,615-618):
// ONE call stages the compressed chunks of all projected columns of all row groups of the batch (H2D), decodes them with
// the fused kernels, evaluates the device filters into a selection mask and copies everything into pinned host buffers.
void FastLanesReader::DecodeRowGroupBatch(FastLanesReadLocalState &ls) {
	const auto &ids = ls.projection.reader_column_ids;
	const auto n_rg = NumericCast<uint32_t>(ls.batch_end - ls.batch_begin);
	const auto rg0 = NumericCast<uint32_t>(ls.batch_begin);
	const idx_t rows = idx_t(n_rg) * info.vec_per_rowgroup * FLS_VEC_SZ;
	while (ls.stage.size() < ids.size()) {
		ls.stage.push_back(make_uniq<FlsColumnStage>());
	}
	// device filters: the selection is applied on the device, only the rows that pass come back (packed, flat)
	const bool select = select_on_device && !ls.device_filters.empty();
	vector<flsgpu_scan_column> scan_columns(ids.size());
	for (idx_t i = 0; i < ids.size(); i++) {
		auto &stage = *ls.stage[i];
		auto &sc = scan_columns[i];
		sc = flsgpu_scan_column {};
		sc.col = ids[i];
		sc.out_type = column_out_types[ids[i]];
		const auto physical = columns[ids[i]].type.InternalType();
		stage.is_string = physical == PhysicalType::VARCHAR;
		stage.dictionary_form = false;
		stage.dictionaries.clear();
		if (stage.is_string) {
			// dictionary-encoded chunks are handed to DuckDB as DICTIONARY_VECTORs like the reference does
			// (dictionary_kernel.hpp:80-101, fsst_dict_kernel.hpp:54-80); taken when every chunk of the batch is
			// dictionary encoded, else the batch is materialised flat
			uint64_t n_keys = 0, heap_bytes = 0;
			stage.dictionary_form = flsgpu_string_dict_keys(table, ids[i], rg0, n_rg, &n_keys) == FLSGPU_OK;
			FlsCheck(flsgpu_string_heap_bytes(table, ids[i], rg0, n_rg, &heap_bytes), "flsgpu_string_heap_bytes");
			stage.value_bytes = stage.dictionary_form ? sizeof(sel_t) : sizeof(string_t);
			sc.string_form = stage.dictionary_form ? FLSGPU_STRING_DICTIONARY : FLSGPU_STRING_FLAT;
			ReservePinned(stage.heap, heap_bytes + 16);
			sc.heap_host = stage.heap->ptr;
			sc.heap_capacity_bytes = stage.heap->bytes;
			if (stage.dictionary_form) {
				ReservePinned(stage.keys, n_keys * sizeof(string_t) + 16);
				sc.keys_host = stage.keys->ptr;
				sc.keys_capacity_bytes = stage.keys->bytes;
				stage.key_offsets.assign(n_rg + 1, 0);
				for (uint32_t r = 0; r < n_rg; r++) {
					fls_chunk_desc chunk;
					FlsCheck(flsgpu_table_chunk(table, rg0 + r, ids[i], &chunk), "flsgpu_table_chunk");
					stage.key_offsets[r + 1] = stage.key_offsets[r] + chunk.aux_count;
				}
				// packed batches (select): one dictionary over the keys of all row groups of the batch, the entries index it
				stage.dictionaries.resize(select ? 1 : n_rg);
				stage.dictionary_ids.assign(select ? 1 : n_rg, string());
			}
		} else {
			stage.value_bytes = GetTypeIdSize(physical);
		}
		stage.forms.clear();
		idx_t stage_rows = rows;
		if (!stage.is_string && compressed_vectors && !select) {
			// vectors the writer marked CONSTANT or arithmetic SEQUENCE are handed to DuckDB in that form (the reference
			// does so for dec_constant_opr: constant_kernel.hpp:19-27): only the FLAT vectors are decoded and cross PCIe
			uint64_t n_flat = 0;
			stage.forms.resize(idx_t(n_rg) * info.vec_per_rowgroup);
			FlsCheck(flsgpu_column_vector_forms(table, ids[i], rg0, n_rg, stage.forms.data(), &n_flat), "flsgpu_column_vector_forms");
			sc.compact = 1;
			sc.forms = stage.forms.data();
			stage_rows = n_flat * FLS_VEC_SZ;
		}
		ReservePinned(stage.rows, stage_rows * stage.value_bytes);
		sc.out_host = stage.rows->ptr;
		sc.out_capacity_bytes = stage.rows->bytes;
	}
	ls.compacted = select;
	ls.n_selected = ls.sel_cursor = 0;
	if (select) {
		bool need_row_index = false;
		for (idx_t i = 0; i < column_ids.size(); i++) {
			need_row_index |= IsFileRowNumberColumn(column_ids[MultiFileLocalIndex(i)].GetId());
		}
		if (need_row_index) {
			ReservePinned(ls.row_index, rows * sizeof(uint32_t));
		}
		uint64_t n_selected = 0;
		FlsCheck(flsgpu_scan_selected_to_host(device_tables[ls.device_slot], scan_columns.data(), NumericCast<uint32_t>(scan_columns.size()),
		                                      ls.device_filters.data(), NumericCast<uint32_t>(ls.device_filters.size()), rg0, n_rg,
		                                      need_row_index ? reinterpret_cast<uint32_t *>(ls.row_index->ptr) : nullptr,
		                                      need_row_index ? ls.row_index->bytes : 0, &n_selected),
		         "read_fls");
		ls.n_selected = n_selected;
		ls.has_mask = false;
		ls.batch_decoded = true;
		return;
	}
	ls.has_mask = !ls.device_filters.empty();
	if (ls.has_mask) {
		ReservePinned(ls.mask, idx_t(n_rg) * info.vec_per_rowgroup * 128);
	}
	if (!scan_columns.empty() || ls.has_mask) {
		FlsCheck(flsgpu_scan_filtered_to_host(device_tables[ls.device_slot], scan_columns.data(), NumericCast<uint32_t>(scan_columns.size()),
		                                      ls.device_filters.data(), NumericCast<uint32_t>(ls.device_filters.size()), rg0, n_rg,
		                                      ls.has_mask ? reinterpret_cast<uint32_t *>(ls.mask->ptr) : nullptr,
		                                      ls.has_mask ? ls.mask->bytes : 0),
		         "read_fls");
	}
	ls.batch_decoded = true;
}

// FastLanesReader::GetScanBatch (fls_reader.cpp:422-434): at most two FastLanes vectors = one DuckDB chunk, never across a
// row group; the partial tail of a row group is exposed as-is
FastLanesReader::ScanBatch FastLanesReader::GetScanBatch(const FastLanesReadLocalState &local_state) const {
	ScanBatch batch;
	batch.start_tuple = local_state.cur_vector * FLS_VEC_SZ;
	if (batch.start_tuple >= local_state.n_tuples) {
		return batch;
	}
	batch.count = MinValue<idx_t>(local_state.n_tuples - batch.start_tuple, STANDARD_VECTOR_SIZE);
	if (batch.count > FLS_VEC_SZ && local_state.batch_decoded &&
	    !VectorsJoin(local_state, (local_state.cur_rowgroup - local_state.batch_begin) * info.vec_per_rowgroup + local_state.cur_vector)) {
		batch.count = FLS_VEC_SZ;
	}
	batch.input_vector_count = (batch.count + FLS_VEC_SZ - 1) / FLS_VEC_SZ;
	return batch;
}

// FastLanesReader::DecodePhysicalColumns (fls_reader.cpp:436-472).  The values are decoded already: every output vector is
// pointed at rows [start_vector * 1024, + count) of the current row group inside the pinned staging buffers — no copy —
// and holds a reference to the buffer (FlatVector data + keep-alive; strings: StringVector::AddBuffer;
// dictionary-encoded strings: Vector::Dictionary over the row group's dictionary, dictionary_kernel.hpp:89-100).
void FastLanesReader::DecodePhysicalColumns(FastLanesReadLocalState &ls, DataChunk &chunk, idx_t start_vector, idx_t count) {
	if (!ls.projection.has_physical_columns) {
		return;
	}
	const idx_t rg_in_batch = ls.cur_rowgroup - ls.batch_begin;
	const idx_t first_row = ls.compacted ? ls.sel_cursor : (rg_in_batch * info.vec_per_rowgroup + start_vector) * FLS_VEC_SZ;
	idx_t n_conversions = 0;
	if (ls.conversion_executor) {
		ls.conversion_input.Reset(); // back to its own buffers: the last chunk left references to the staging memory
	}
	for (idx_t i = 0; i < column_ids.size(); i++) {
		const auto reader_idx = ls.projection.output_to_reader_idx[i];
		if (!reader_idx.IsValid()) {
			continue;
		}
		auto &stage = *ls.stage[reader_idx.GetIndex()];
		const auto &type = columns[ls.projection.reader_column_ids[reader_idx.GetIndex()]].type;
		const bool converted = ls.conversion_of_output[i].IsValid();
		auto &target = converted ? ls.conversion_input.data[n_conversions++] : chunk.data[i];
		if (stage.dictionary_form) {
			const idx_t dict_slot = ls.compacted ? 0 : rg_in_batch;
			const idx_t first_key = ls.compacted ? 0 : stage.key_offsets[rg_in_batch];
			auto &dictionary = stage.dictionaries[dict_slot];
			const idx_t dict_size = (ls.compacted ? stage.key_offsets.back() : stage.key_offsets[rg_in_batch + 1]) - first_key;
			if (!dictionary) { // once per row group, like KernelTraits<dec_dict_opr<fls_string_t>>::Prepare
				dictionary = make_uniq<Vector>(type, stage.keys->ptr + first_key * sizeof(string_t));
				StringVector::AddBuffer(*dictionary, stage.keys);
				StringVector::AddBuffer(*dictionary, stage.heap);
				// operators cache work per dictionary id (hash tables, dictionary-aware functions): it must never repeat,
				// not even when a later dictionary is allocated at the address of a freed one
				static atomic<idx_t> next_dictionary_id {0};
				stage.dictionary_ids[dict_slot] = "fls_dict_" + to_string(next_dictionary_id++);
			}
			// the selection entries of the chunk's rows: copied (8 KiB) so that the dictionary vector owns them
			SelectionVector sel(STANDARD_VECTOR_SIZE);
			std::memcpy(sel.data(), stage.rows->ptr + first_row * sizeof(sel_t), count * sizeof(sel_t));
			target.Dictionary(*dictionary, dict_size, sel, count);
			DictionaryVector::SetDictionaryId(target, stage.dictionary_ids[dict_slot]);
			continue;
		}
		if (!stage.forms.empty()) {
			FormedVector(stage, type, (rg_in_batch * info.vec_per_rowgroup + start_vector), count, target);
			continue;
		}
		Vector view(type, stage.rows->ptr + first_row * stage.value_bytes);
		if (stage.is_string) {
			StringVector::AddBuffer(view, stage.rows);
			StringVector::AddBuffer(view, stage.heap);
		} else {
			view.SetAuxiliary(stage.rows);
		}
		target.Reference(view);
	}
	if (n_conversions) {
		// evaluate the conversion expressions over the decoded columns into the output columns
		ls.conversion_input.SetCardinality(count);
		n_conversions = 0;
		for (idx_t i = 0; i < column_ids.size(); i++) {
			if (ls.conversion_of_output[i].IsValid()) {
				DataChunk one;
				one.InitializeEmpty({ls.conversion_input.data[n_conversions].GetType()});
				one.data[0].Reference(ls.conversion_input.data[n_conversions]);
				one.SetCardinality(count);
				ls.conversion_executor->SetChunk(one);
				ls.conversion_executor->ExecuteExpression(ls.conversion_of_output[i].GetIndex(), chunk.data[i]);
				n_conversions++;
			}
		}
	}
}

template <class T>
static void GenerateSequence(int64_t start, int64_t increment, T *out, idx_t n) {
	uint64_t value = static_cast<uint64_t>(start);
	for (idx_t i = 0; i < n; i++, value += static_cast<uint64_t>(increment)) {
		out[i] = static_cast<T>(value); // exact in the column's type: flsgpu_column_vector_forms guarantees it
	}
}

// Do the two vectors of a would-be chunk continue each other in every compact column?  FLAT + FLAT side by side in the packed
// buffer, or CONSTANT / SEQUENCE halves where the second takes up where the first ends.  If not, the chunk is cut after its
// first vector (1024 rows, like the tail of a row group): every vector stays zero-copy, nothing is put together on the host.
bool FastLanesReader::VectorsJoin(const FastLanesReadLocalState &ls, idx_t first_vector) const {
	for (auto &stage : ls.stage) {
		if (stage->forms.empty()) {
			continue;
		}
		const auto &f0 = stage->forms[first_vector], &f1 = stage->forms[first_vector + 1];
		if (f0.form == FLS_VEC_FORM_FLAT || f1.form == FLS_VEC_FORM_FLAT) {
			if (f0.form != f1.form || f1.flat_slot != f0.flat_slot + 1) {
				return false;
			}
			continue;
		}
		const int64_t inc0 = f0.form == FLS_VEC_FORM_SEQUENCE ? f0.increment : 0, inc1 = f1.form == FLS_VEC_FORM_SEQUENCE ? f1.increment : 0;
		if (inc0 != inc1 || f1.start != f0.start + inc0 * int64_t(FLS_VEC_SZ)) {
			return false;
		}
	}
	return true;
}

// One chunk of a column scanned in compact form (its vectors join, see VectorsJoin): a zero-copy FLAT slice of the packed
// staging buffer, a CONSTANT_VECTOR (what the reference emits for dec_constant_opr, constant_kernel.hpp:19-27) or a
// SEQUENCE_VECTOR; types DuckDB has no sequence vectors for (unsigned, DATE, TIMESTAMP) get the sequence written out.
void FastLanesReader::FormedVector(FlsColumnStage &stage, const LogicalType &type, idx_t first_vector, idx_t count, Vector &target) {
	const auto &f0 = stage.forms[first_vector];
	if (f0.form == FLS_VEC_FORM_FLAT) {
		Vector view(type, stage.rows->ptr + idx_t(f0.flat_slot) * FLS_VEC_SZ * stage.value_bytes);
		view.SetAuxiliary(stage.rows);
		target.Reference(view);
		return;
	}
	const auto physical = type.InternalType();
	if (f0.form == FLS_VEC_FORM_CONSTANT) {
		target.SetVectorType(VectorType::CONSTANT_VECTOR);
		ConstantVector::SetNull(target, false);
		std::memcpy(ConstantVector::GetData(target), &f0.start, stage.value_bytes); // little endian: the low bytes are the value
		return;
	}
	// DuckDB generates sequences for signed integer vectors only (generators.cpp:30-50)
	if (type.IsNumeric() && (physical == PhysicalType::INT8 || physical == PhysicalType::INT16 || physical == PhysicalType::INT32 ||
	                         physical == PhysicalType::INT64)) {
		target.Sequence(f0.start, f0.increment, count);
		return;
	}
	target.SetVectorType(VectorType::FLAT_VECTOR);
	const auto out = FlatVector::GetData(target);
	switch (stage.value_bytes) {
	case 1:
		GenerateSequence(f0.start, f0.increment, reinterpret_cast<uint8_t *>(out), count);
		break;
	case 2:
		GenerateSequence(f0.start, f0.increment, reinterpret_cast<uint16_t *>(out), count);
		break;
	case 4:
		GenerateSequence(f0.start, f0.increment, reinterpret_cast<uint32_t *>(out), count);
		break;
	default:
		GenerateSequence(f0.start, f0.increment, reinterpret_cast<uint64_t *>(out), count);
		break;
	}
}

// FastLanesReader::PopulateVirtualColumns (fls_reader.cpp:474-495): file_row_number = row_group_base + start + i
void FastLanesReader::PopulateVirtualColumns(const FastLanesReadLocalState &ls, DataChunk &chunk, idx_t row_start, idx_t count) {
	for (idx_t i = 0; i < column_ids.size(); i++) {
		if (!IsFileRowNumberColumn(column_ids[MultiFileLocalIndex(i)].GetId())) {
			continue;
		}
		auto &target = chunk.data[i];
		target.SetVectorType(VectorType::FLAT_VECTOR);
		const auto data = FlatVector::GetData<int64_t>(target);
		if (ls.compacted) { // packed rows: their positions inside the batch came back with them
			const auto positions = reinterpret_cast<const uint32_t *>(ls.row_index->ptr) + ls.sel_cursor;
			const idx_t rg_rows = idx_t(info.vec_per_rowgroup) * FLS_VEC_SZ;
			for (idx_t row = 0; row < count; row++) {
				data[row] = NumericCast<int64_t>(row_group_offsets[ls.batch_begin + positions[row] / rg_rows] + positions[row] % rg_rows);
			}
			continue;
		}
		for (idx_t row = 0; row < count; row++) {
			data[row] = NumericCast<int64_t>(row_start + row);
		}
	}
}

// FilterExecutor::Apply (filter_executor.cpp:19-60).  The device already evaluated its share of the filters for the whole
// batch (one bit per row); the mask words of this chunk become the selection vector, the host filters narrow it further
// with ColumnSegment::FilterSelection, exactly the call the reference makes for every filter.
void FastLanesReader::ApplyFilters(FastLanesReadLocalState &ls, DataChunk &chunk, idx_t start_vector, idx_t count) {
	if (!ls.has_mask && ls.host_filters.empty()) {
		return;
	}
	// a fresh selection vector per chunk: the chunk's vectors share its buffer after Slice, and DuckDB's filter functions
	// replace the vector they are handed by a smaller one (column_segment.cpp: sel.Initialize(new_sel))
	SelectionVector selection(STANDARD_VECTOR_SIZE);
	idx_t active = 0;
	auto sel = selection.data();
	if (ls.has_mask) {
		const idx_t rg_in_batch = ls.cur_rowgroup - ls.batch_begin;
		const auto words = reinterpret_cast<const uint32_t *>(ls.mask->ptr) + (rg_in_batch * info.vec_per_rowgroup + start_vector) * 32;
		for (idx_t w = 0; w < (count + 31) / 32; w++) {
			uint32_t bits = words[w];
			if (w * 32 + 32 > count) {
				bits &= count % 32 ? (1u << (count % 32)) - 1 : ~0u;
			}
			while (bits) {
				sel[active++] = NumericCast<sel_t>(w * 32 + NumericCast<idx_t>(__builtin_ctz(bits)));
				bits &= bits - 1;
			}
		}
	} else {
		for (idx_t i = 0; i < count; i++) {
			sel[i] = NumericCast<sel_t>(i);
		}
		active = count;
	}
	for (auto &host_filter : ls.host_filters) {
		if (active == 0) {
			break;
		}
		auto &vector = chunk.data[host_filter.chunk_column];
		UnifiedVectorFormat unified;
		vector.ToUnifiedFormat(count, unified);
		ColumnSegment::FilterSelection(selection, vector, unified, host_filter.filter, *host_filter.filter_state, count, active);
	}
	if (active != count) {
		chunk.Slice(selection, active);
	}
}

// FastLanesReader::Scan (fls_reader.cpp:516-547).  Leaves the chunk empty when the batch is exhausted; MultiFileScan then
// asks TryInitializeScan for the next batch (multi_file_function.hpp:611).
void FastLanesReader::Scan(ClientContext &, GlobalTableFunctionState &, LocalTableFunctionState &local_state_p, DataChunk &chunk) {
	auto &ls = local_state_p.Cast<FastLanesReadLocalState>();
	if (ls.batch_begin >= ls.batch_end) {
		return;
	}
	if (!ls.batch_decoded) {
		DecodeRowGroupBatch(ls);
		if (ls.conversion_executor && ls.conversion_input.ColumnCount() == 0) {
			vector<LogicalType> input_types;
			for (idx_t i = 0; i < column_ids.size(); i++) {
				if (ls.conversion_of_output[i].IsValid()) {
					input_types.push_back(columns[column_ids[MultiFileLocalIndex(i)].GetId()].type);
				}
			}
			ls.conversion_input.Initialize(Allocator::DefaultAllocator(), input_types);
		}
	}
	while (ls.compacted) { // the batch is a packed run of selected rows: hand out 2048-row slices of it
		if (ls.sel_cursor >= ls.n_selected) {
			for (idx_t r = ls.batch_begin; r < ls.batch_end; r++) {
				vectors_read += GetNVectors(r);
			}
			ls.batch_begin = ls.batch_end; // batch exhausted
			return;
		}
		const idx_t count = MinValue<idx_t>(STANDARD_VECTOR_SIZE, ls.n_selected - ls.sel_cursor);
		DecodePhysicalColumns(ls, chunk, 0, count);
		PopulateVirtualColumns(ls, chunk, 0, count);
		chunk.SetCardinality(count);
		ApplyFilters(ls, chunk, 0, count); // what the device did not evaluate
		ls.sel_cursor += count;
		if (chunk.size() == 0) {
			chunk.Reset();
			continue;
		}
		return;
	}
	while (true) {
		const auto batch = GetScanBatch(ls);
		if (!batch.IsValid()) {
			if (ls.cur_rowgroup + 1 >= ls.batch_end) {
				ls.batch_begin = ls.batch_end; // batch exhausted
				return;
			}
			ls.cur_rowgroup++;
			ls.row_group_base = row_group_offsets[ls.cur_rowgroup];
			ls.n_tuples = GetNTuples(ls.cur_rowgroup);
			ls.cur_vector = 0;
			continue;
		}
		DecodePhysicalColumns(ls, chunk, ls.cur_vector, batch.count);
		PopulateVirtualColumns(ls, chunk, ls.row_group_base + batch.start_tuple, batch.count);
		chunk.SetCardinality(batch.count);
		ApplyFilters(ls, chunk, ls.cur_vector, batch.count);
		ls.cur_vector += batch.input_vector_count;
		vectors_read += batch.input_vector_count;
		if (chunk.size() == 0) {
			chunk.Reset();
			continue; // every row of the chunk was filtered out
		}
		return;
	}
}

} // namespace duckdb
