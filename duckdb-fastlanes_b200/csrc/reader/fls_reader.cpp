// fls_reader.cpp — see fls_reader.hpp.  Every method cites the reference method it stands in for.
#include "fls_reader.hpp"

#include <algorithm>
#include <cstring>

namespace flsreader {

namespace {
void Check(int rc, const char* where) {
	if (rc != FLSGPU_OK) {
		throw ReaderError(rc, std::string(where) + ": " + flsgpu_last_error());
	}
}
} // namespace

PinnedBuffer::~PinnedBuffer() {
	if (ptr) {
		flsgpu_host_free(ptr);
	}
}
void PinnedBuffer::Reserve(uint64_t n) {
	if (n <= bytes) {
		return;
	}
	if (ptr) {
		flsgpu_host_free(ptr);
		ptr = nullptr;
		bytes = 0;
	}
	Check(flsgpu_host_alloc(n, &ptr), "flsgpu_host_alloc");
	bytes = n;
}

// TranslateUtils::TranslateType (reference src/reader/translation_utils.cpp:5-48).  The out_type is the DuckDB
// physical type the materializer writes for that logical type (fls_type_resolver.hpp:42-83): every unsigned
// width is widened to UBIGINT (:19-24), everything else keeps its physical width.
LogicalType TranslateType(const fls_column_desc& col, int32_t* out_type) {
	LogicalType t;
	int32_t ot = FLSGPU_OUT_NATIVE;
	switch (col.data_type) {
	case FLS_TYPE_INT8: t.id = LogicalTypeId::TINYINT; t.physical_bytes = 1; break;
	case FLS_TYPE_INT16: t.id = LogicalTypeId::SMALLINT; t.physical_bytes = 2; break;
	case FLS_TYPE_INT32: t.id = LogicalTypeId::INTEGER; t.physical_bytes = 4; break;
	case FLS_TYPE_INT64: t.id = LogicalTypeId::BIGINT; t.physical_bytes = 8; break;
	case FLS_TYPE_UINT8:
	case FLS_TYPE_UINT16:
	case FLS_TYPE_UINT32:
	case FLS_TYPE_UINT64: t.id = LogicalTypeId::UBIGINT; t.physical_bytes = 8; ot = FLSGPU_OUT_U64; break;
	case FLS_TYPE_FLOAT: t.id = LogicalTypeId::FLOAT; t.physical_bytes = 4; break;
	case FLS_TYPE_DOUBLE: t.id = LogicalTypeId::DOUBLE; t.physical_bytes = 8; break;
	case FLS_TYPE_BOOLEAN: t.id = LogicalTypeId::BOOLEAN; t.physical_bytes = 1; break;
	case FLS_TYPE_DATE: t.id = LogicalTypeId::DATE; t.physical_bytes = 4; break;
	case FLS_TYPE_TIMESTAMP: t.id = LogicalTypeId::TIMESTAMP; t.physical_bytes = 8; break;
	case FLS_TYPE_DECIMAL: // fls_reader.cpp:119-126: DECIMAL(width, scale) from the column annotation
		t.id = LogicalTypeId::DECIMAL;
		t.width = col.decimal_precision;
		t.scale = col.decimal_scale;
		t.physical_bytes = col.phys_bits / 8;
		break;
	default: throw ReaderError(FLSGPU_ERR_UNSUPPORTED, "TranslateType: unsupported FastLanes data type"); // :44-46
	}
	if (out_type) {
		*out_type = ot;
	}
	return t;
}

FastLanesReader::FastLanesReader(const std::string& path, const FastLanesFileReaderOptions& options) : reader_options(options) {
	Check(flsgpu_table_open(path.c_str(), &table), "flsgpu_table_open"); // TableMetadata (table_metadata.cpp:13-16)
	try {
		Initialize();
	} catch (...) {
		flsgpu_table_close(table);
		table = nullptr;
		throw;
	}
}

FastLanesReader::~FastLanesReader() {
	flsgpu_table_close(table);
}

// FastLanesReader::Initialize (fls_reader.cpp:106-143): schema, per-row-group tuple offsets, totals.
void FastLanesReader::Initialize() {
	Check(flsgpu_table_info_get(table, &info), "flsgpu_table_info_get");
	columns.clear();
	column_out_types.clear();
	for (uint32_t c = 0; c < info.n_columns; ++c) {
		fls_column_desc d;
		Check(flsgpu_table_column(table, c, &d), "flsgpu_table_column");
		int32_t ot = 0;
		ColumnDefinition def;
		def.name = std::string(d.name, strnlen(d.name, sizeof(d.name)));
		def.type = TranslateType(d, &ot);
		columns.push_back(def);
		column_out_types.push_back(ot);
	}
	row_groups.resize(info.n_rowgroups);
	row_group_offsets.resize(info.n_rowgroups);
	total_tuples = 0;
	total_vectors = 0;
	for (uint32_t r = 0; r < info.n_rowgroups; ++r) {
		Check(flsgpu_table_rowgroup(table, r, &row_groups[r]), "flsgpu_table_rowgroup");
		row_group_offsets[r] = total_tuples;
		total_tuples += row_groups[r].n_tuples;
		total_vectors += row_groups[r].n_vec;
	}
	if (reader_options.file_row_number) { // EnsureFileRowNumberColumn
		ColumnDefinition def;
		def.name = "file_row_number";
		def.type.id = LogicalTypeId::BIGINT;
		def.type.physical_bytes = 8;
		columns.push_back(def);
	}
	if (reader_options.rowgroups_per_batch == 0) {
		reader_options.rowgroups_per_batch = 1;
	}
}

idx_t FastLanesReader::GetNRowGroups() const {
	return info.n_rowgroups;
}
idx_t FastLanesReader::GetNTuples(idx_t row_group_idx) const {
	return row_groups.at(row_group_idx).n_tuples;
}
idx_t FastLanesReader::GetNVectors(idx_t row_group_idx) const {
	return row_groups.at(row_group_idx).n_vec;
}
bool FastLanesReader::IsFileRowNumberColumn(idx_t column_id) const {
	return column_id == COLUMN_IDENTIFIER_FILE_ROW_NUMBER ||
	       (reader_options.file_row_number && column_id == (idx_t)info.n_columns);
}

// FastLanesMultiFileInfo::GetCardinality (fls_multi_file_info.cpp:152-164)
idx_t FastLanesReader::GetCardinality() const {
	return reader_options.has_explicit_cardinality ? reader_options.explicit_cardinality : total_tuples;
}
// FastLanesMultiFileInfo::MaxThreads (fls_multi_file_info.cpp:99-110): one worker per unit of hand-out
idx_t FastLanesReader::MaxThreads() const {
	return std::max<idx_t>(1, (info.n_rowgroups + reader_options.rowgroups_per_batch - 1) / reader_options.rowgroups_per_batch);
}

// FastLanesReader::BuildReaderProjectionPlan (fls_reader.cpp:560-613).  The multi-column-compression fallback
// (HasMccEncoding, :43-82) has nothing to do here: this container has no cross-column encodings.
FastLanesReaderProjectionPlan FastLanesReader::BuildReaderProjectionPlan() const {
	FastLanesReaderProjectionPlan plan;
	plan.output_to_reader_idx.assign(column_ids.size(), -1);
	for (size_t i = 0; i < column_ids.size(); ++i) {
		const idx_t id = column_ids[i];
		if (IsFileRowNumberColumn(id)) {
			continue;
		}
		if (id >= info.n_columns) {
			throw ReaderError(FLSGPU_ERR_INVALID, "Projected column id out of range"); // reference :577 InternalException
		}
		auto it = std::find(plan.reader_column_ids.begin(), plan.reader_column_ids.end(), (uint32_t)id);
		if (it == plan.reader_column_ids.end()) {
			plan.output_to_reader_idx[i] = (int64_t)plan.reader_column_ids.size();
			plan.reader_column_ids.push_back((uint32_t)id);
		} else {
			plan.output_to_reader_idx[i] = it - plan.reader_column_ids.begin();
		}
	}
	plan.has_physical_columns = !plan.reader_column_ids.empty();
	return plan;
}

// FastLanesReader::TryAssignNextRowGroup (fls_reader.cpp:337-355), handing out a batch of row groups.
bool FastLanesReader::TryAssignNextRowGroup(FastLanesReadGlobalState& global_state, FastLanesReadLocalState& local_state) {
	if (global_state.next_rowgroup >= info.n_rowgroups) {
		return false;
	}
	local_state.batch_begin = global_state.next_rowgroup;
	local_state.batch_end = std::min<idx_t>(info.n_rowgroups, local_state.batch_begin + reader_options.rowgroups_per_batch);
	local_state.cur_rowgroup = local_state.batch_begin;
	local_state.projection = BuildReaderProjectionPlan();
	local_state.row_group_base = row_group_offsets[local_state.cur_rowgroup];
	local_state.cur_vector = 0;
	local_state.n_tuples = GetNTuples(local_state.cur_rowgroup);
	return true;
}

// FastLanesReader::TryInitializeScan (fls_reader.cpp:497-514).  The caller holds gstate.lock, exactly like DuckDB's
// MultiFileGlobalState::lock (multi_file_function.hpp:399-416); the GPU work happens outside the lock in the first Scan.
bool FastLanesReader::TryInitializeScan(FastLanesReadGlobalState& gstate, FastLanesReadLocalState& lstate) {
	if (!TryAssignNextRowGroup(gstate, lstate)) {
		return false;
	}
	gstate.next_rowgroup = lstate.batch_end; // mark the row groups as consumed (:511)
	lstate.staging.resize(lstate.projection.reader_column_ids.size());
	lstate.row_numbers.clear();
	lstate.cur_vector = ~0ull; // batch not decoded yet
	return true;
}

// get_rowgroup_reader + InitializeColumnDecoders + every Decode<PASS> of the batch (fls_reader.cpp:382-420,436-472):
// one pipelined H2D -> decode -> D2H call for all projected columns of all row groups of the batch.
void FastLanesReader::DecodeRowGroupBatch(FastLanesReadLocalState& ls) {
	const auto& ids = ls.projection.reader_column_ids;
	const uint32_t n_rg = (uint32_t)(ls.batch_end - ls.batch_begin);
	if (!ids.empty()) {
		std::vector<int32_t> out_types(ids.size());
		std::vector<void*> ptrs(ids.size());
		std::vector<uint64_t> caps(ids.size());
		for (size_t i = 0; i < ids.size(); ++i) {
			out_types[i] = column_out_types[ids[i]];
			const uint64_t need = (uint64_t)n_rg * info.vec_per_rowgroup * FLS_VEC_SZ * columns[ids[i]].type.physical_bytes;
			ls.staging[i].Reserve(need);
			ptrs[i] = ls.staging[i].ptr;
			caps[i] = ls.staging[i].bytes;
		}
		Check(flsgpu_scan_columns_to_host(table, ids.data(), (uint32_t)ids.size(), (uint32_t)ls.batch_begin, n_rg, out_types.data(),
		                                  ptrs.data(), caps.data()),
		      "flsgpu_scan_columns_to_host");
	}
	ls.cur_vector = 0;
}

// FastLanesReader::GetScanBatch (fls_reader.cpp:422-434): at most two FastLanes vectors = one DuckDB chunk, never
// across a row group, partial tail exposed as-is.
FastLanesReader::ScanBatch FastLanesReader::GetScanBatch(const FastLanesReadLocalState& local_state) const {
	ScanBatch batch;
	batch.start_tuple = local_state.cur_vector * FLS_VEC_SZ;
	if (batch.start_tuple >= local_state.n_tuples) {
		return batch;
	}
	const idx_t tuples_left = local_state.n_tuples - batch.start_tuple;
	batch.count = std::min<idx_t>(tuples_left, FLS_VEC_SZ * 2);
	batch.input_vector_count = (batch.count + FLS_VEC_SZ - 1) / FLS_VEC_SZ;
	return batch;
}

// FastLanesReader::DecodePhysicalColumns (fls_reader.cpp:436-472): the values are already decoded; point the chunk's
// vectors at rows [start_vector*1024, +count) of the current row group inside the staging buffers.
void FastLanesReader::DecodePhysicalColumns(FastLanesReadLocalState& ls, DataChunk& chunk, idx_t start_vector, idx_t) {
	const idx_t rg_in_batch = ls.cur_rowgroup - ls.batch_begin;
	const idx_t first_elem = (rg_in_batch * info.vec_per_rowgroup + start_vector) * FLS_VEC_SZ;
	for (size_t i = 0; i < column_ids.size(); ++i) {
		const int64_t ri = ls.projection.output_to_reader_idx[i];
		if (ri < 0) {
			continue;
		}
		const ColumnDefinition& def = columns[ls.projection.reader_column_ids[ri]];
		chunk.data[i].type = def.type;
		chunk.data[i].data = static_cast<const uint8_t*>(ls.staging[ri].ptr) + first_elem * def.type.physical_bytes;
	}
}

// FastLanesReader::PopulateVirtualColumns (fls_reader.cpp:474-495): file_row_number = row_group_base + start + i
void FastLanesReader::PopulateVirtualColumns(FastLanesReadLocalState& ls, DataChunk& chunk, idx_t row_start, idx_t count) {
	for (size_t i = 0; i < column_ids.size(); ++i) {
		if (!IsFileRowNumberColumn(column_ids[i])) {
			continue;
		}
		ls.row_numbers.resize(STANDARD_VECTOR_SIZE);
		for (idx_t row = 0; row < count; ++row) {
			ls.row_numbers[row] = (int64_t)(row_start + row);
		}
		chunk.data[i].type.id = LogicalTypeId::BIGINT;
		chunk.data[i].type.physical_bytes = 8;
		chunk.data[i].data = ls.row_numbers.data();
	}
}

// FastLanesReader::Scan (fls_reader.cpp:516-547).  Returns with chunk.count == 0 when the batch is exhausted; the
// caller then asks TryInitializeScan for the next one (multi_file_function.hpp:611).
void FastLanesReader::Scan(FastLanesReadGlobalState&, FastLanesReadLocalState& ls, DataChunk& chunk) {
	chunk.data.resize(column_ids.size());
	chunk.SetCardinality(0);
	if (ls.batch_begin >= ls.batch_end) {
		return;
	}
	if (ls.cur_vector == ~0ull) {
		DecodeRowGroupBatch(ls);
	}
	while (true) {
		const ScanBatch batch = GetScanBatch(ls);
		if (!batch.IsValid()) {
			// next row group of the batch
			if (ls.cur_rowgroup + 1 >= ls.batch_end) {
				ls.batch_begin = ls.batch_end; // exhausted
				return;
			}
			ls.cur_rowgroup++;
			ls.row_group_base = row_group_offsets[ls.cur_rowgroup];
			ls.n_tuples = GetNTuples(ls.cur_rowgroup);
			ls.cur_vector = 0;
			continue;
		}
		DecodePhysicalColumns(ls, chunk, ls.cur_vector, batch.input_vector_count);
		const idx_t row_start = ls.row_group_base + batch.start_tuple;
		PopulateVirtualColumns(ls, chunk, row_start, batch.count);
		chunk.SetCardinality(batch.count);
		ls.cur_vector += batch.input_vector_count;
		vectors_read += batch.input_vector_count;
		return;
	}
}

// FastLanesReader::GetProgressInFile (fls_reader.cpp:556-558)
double FastLanesReader::GetProgressInFile() const {
	return total_vectors ? 100.0 * ((double)vectors_read.load() / (double)total_vectors) : 100.0;
}

// FastLanesReader::GetPartitionStats (fls_reader.cpp:316-332): exact row ranges per row group (COUNT(*) pushdown)
void FastLanesReader::GetPartitionStats(std::vector<PartitionStatistics>& result) const {
	for (uint32_t r = 0; r < info.n_rowgroups; ++r) {
		PartitionStatistics p;
		p.row_start = row_group_offsets[r];
		p.count = row_groups[r].n_tuples;
		result.push_back(p);
	}
}

// FastLanesReader::GetStatistics (fls_reader.cpp:244-292): file-level min/max merged from the per-chunk zone maps
// (RowGroupStatistics, row_group_statistics.cpp:60-127; numeric columns only).
bool FastLanesReader::GetStatistics(const std::string& name, NumericStats& out) const {
	for (uint32_t c = 0; c < info.n_columns; ++c) {
		if (columns[c].name != name) {
			continue;
		}
		fls_column_desc cd;
		Check(flsgpu_table_column(table, c, &cd), "flsgpu_table_column");
		out = NumericStats();
		for (uint32_t r = 0; r < info.n_rowgroups; ++r) {
			fls_chunk_desc ch;
			Check(flsgpu_table_chunk(table, r, c, &ch), "flsgpu_table_chunk");
			if (!ch.has_stats) {
				return false;
			}
			if (cd.is_float) {
				double mn, mx;
				std::memcpy(&mn, &ch.min_val, 8);
				std::memcpy(&mx, &ch.max_val, 8);
				out.min_d = out.has_stats ? std::min(out.min_d, mn) : mn;
				out.max_d = out.has_stats ? std::max(out.max_d, mx) : mx;
			} else if (cd.is_signed) {
				out.min_i = out.has_stats ? std::min(out.min_i, (int64_t)ch.min_val) : (int64_t)ch.min_val;
				out.max_i = out.has_stats ? std::max(out.max_i, (int64_t)ch.max_val) : (int64_t)ch.max_val;
			} else {
				const uint64_t mn = out.has_stats ? std::min((uint64_t)out.min_i, ch.min_val) : ch.min_val;
				const uint64_t mx = out.has_stats ? std::max((uint64_t)out.max_i, ch.max_val) : ch.max_val;
				out.min_i = (int64_t)mn;
				out.max_i = (int64_t)mx;
			}
			out.has_stats = true;
		}
		return out.has_stats;
	}
	return false;
}

} // namespace flsreader
