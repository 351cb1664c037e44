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
	case FLS_TYPE_FLS_STR: t.id = LogicalTypeId::VARCHAR; t.physical_bytes = 16; break; // :38-39 (string_t)
	case FLS_TYPE_BYTE_ARRAY: t.id = LogicalTypeId::BLOB; t.physical_bytes = 16; break;
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
	for (size_t i = 1; i < device_tables.size(); ++i) {
		flsgpu_table_close(device_tables[i]);
	}
	flsgpu_table_close(table);
}

// FastLanesReader::Initialize (fls_reader.cpp:106-143): schema, per-row-group tuple offsets, totals.
void FastLanesReader::Initialize() {
	Check(flsgpu_table_info_get(table, &info), "flsgpu_table_info_get");
	int n_devices = 1;
	Check(flsgpu_n_devices(&n_devices), "flsgpu_n_devices");
	device_tables.assign(1, table);
	for (int g = 1; g < n_devices; ++g) {
		flsgpu_table* shared = nullptr;
		Check(flsgpu_table_share(table, &shared), "flsgpu_table_share");
		device_tables.push_back(shared);
		Check(flsgpu_table_bind_device(shared, g), "flsgpu_table_bind_device");
	}
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
// (HasMccEncoding, :43-82: the reference decodes EVERY column of a row group when a projected one is EXP_EQUAL / borrows a
// dictionary) is not needed: an EXP_EQUAL chunk of this container carries a copy of its source chunk's descriptor (same
// offset and segments, fls_chunk_desc.is_equal), so it stages and decodes like any other chunk, projected alone or not.
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
	// RowGroupFilter: row groups whose zone maps rule the filters out are never handed out (row_group_filter.cpp:80-94)
	if (!filters.empty() && !rowgroup_filter_built) {
		BuildRowGroupFilter();
	}
	const bool pruning = !filters.empty();
	while (pruning && global_state.next_rowgroup < info.n_rowgroups && !rowgroups_to_scan[global_state.next_rowgroup]) {
		global_state.next_rowgroup++;
	}
	if (global_state.next_rowgroup >= info.n_rowgroups) {
		return false;
	}
	local_state.batch_begin = global_state.next_rowgroup;
	local_state.batch_end = std::min<idx_t>(info.n_rowgroups, local_state.batch_begin + reader_options.rowgroups_per_batch);
	for (idx_t r = local_state.batch_begin; pruning && r < local_state.batch_end; ++r) { // a batch is a run of kept row groups
		if (!rowgroups_to_scan[r]) {
			local_state.batch_end = r;
		}
	}
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
	lstate.device_slot = gstate.next_batch++ % device_tables.size();
	const size_t n_reader_cols = lstate.projection.reader_column_ids.size();
	lstate.staging.resize(n_reader_cols);
	lstate.heaps.resize(n_reader_cols);
	lstate.keys.resize(n_reader_cols);
	lstate.dict_form.assign(n_reader_cols, 0);
	lstate.forms.assign(n_reader_cols, {});
	lstate.key_offsets.assign(n_reader_cols, {});
	lstate.has_mask = false;
	lstate.row_numbers.clear();
	lstate.cur_vector = ~0ull; // batch not decoded yet
	return true;
}

// get_rowgroup_reader + InitializeColumnDecoders + every Decode<PASS> of the batch (fls_reader.cpp:382-420,436-472):
// one pipelined H2D -> decode -> D2H call for all projected columns of all row groups of the batch.
void FastLanesReader::DecodeRowGroupBatch(FastLanesReadLocalState& ls) {
	const auto& ids = ls.projection.reader_column_ids;
	const uint32_t n_rg = (uint32_t)(ls.batch_end - ls.batch_begin), rg0 = (uint32_t)ls.batch_begin;
	const std::vector<flsgpu_filter> phys_filters = PhysicalFilters();
	if (!ids.empty()) {
		std::vector<flsgpu_scan_column> sc(ids.size());
		for (size_t i = 0; i < ids.size(); ++i) {
			const LogicalType& type = columns[ids[i]].type;
			flsgpu_scan_column& c = sc[i];
			c = flsgpu_scan_column {};
			c.col = ids[i];
			c.out_type = column_out_types[ids[i]];
			const uint64_t rows = (uint64_t)n_rg * info.vec_per_rowgroup * FLS_VEC_SZ;
			if (type.id == LogicalTypeId::VARCHAR || type.id == LogicalTypeId::BLOB) {
				// dec_dict_opr<fls_string_t> / dec_fsst_dict_opr hand DuckDB a DICTIONARY_VECTOR (dictionary_kernel.hpp:80-101);
				// taken here when every chunk of the batch is dictionary encoded, else the batch is materialised flat
				uint64_t n_keys = 0, heap_bytes = 0;
				const bool dict = flsgpu_string_dict_keys(table, ids[i], rg0, n_rg, &n_keys) == FLSGPU_OK;
				Check(flsgpu_string_heap_bytes(table, ids[i], rg0, n_rg, &heap_bytes), "flsgpu_string_heap_bytes");
				ls.dict_form[i] = dict ? 1 : 0;
				c.string_form = dict ? FLSGPU_STRING_DICTIONARY : FLSGPU_STRING_FLAT;
				ls.staging[i].Reserve(rows * (dict ? 4u : 16u));
				ls.heaps[i].Reserve(heap_bytes + 16);
				c.heap_host = ls.heaps[i].ptr;
				c.heap_capacity_bytes = ls.heaps[i].bytes;
				if (dict) {
					ls.keys[i].Reserve(n_keys * 16 + 16);
					c.keys_host = ls.keys[i].ptr;
					c.keys_capacity_bytes = ls.keys[i].bytes;
					ls.key_offsets[i].assign(n_rg + 1, 0);
					for (uint32_t r = 0; r < n_rg; ++r) {
						fls_chunk_desc ch;
						Check(flsgpu_table_chunk(table, rg0 + r, ids[i], &ch), "flsgpu_table_chunk");
						ls.key_offsets[i][r + 1] = ls.key_offsets[i][r] + ch.aux_count;
					}
				}
			} else if (reader_options.compressed_vectors) {
				// only the FLAT vectors of the batch are decoded and cross PCIe, packed (flsgpu_scan_column.compact)
				uint64_t n_flat = 0;
				ls.forms[i].resize((size_t)n_rg * info.vec_per_rowgroup);
				Check(flsgpu_column_vector_forms(table, ids[i], rg0, n_rg, ls.forms[i].data(), &n_flat), "flsgpu_column_vector_forms");
				c.compact = 1;
				c.forms = ls.forms[i].data();
				ls.staging[i].Reserve(std::max<uint64_t>(16, n_flat * FLS_VEC_SZ * type.physical_bytes));
			} else {
				ls.staging[i].Reserve(rows * type.physical_bytes);
			}
			c.out_host = ls.staging[i].ptr;
			c.out_capacity_bytes = ls.staging[i].bytes;
		}
		if (!phys_filters.empty()) {
			ls.mask.Reserve((uint64_t)n_rg * info.vec_per_rowgroup * 128u);
			ls.has_mask = true;
		}
		Check(flsgpu_scan_filtered_to_host(device_tables[ls.device_slot], sc.data(), (uint32_t)sc.size(), phys_filters.data(), (uint32_t)phys_filters.size(), rg0, n_rg,
		                                   static_cast<uint32_t*>(ls.mask.ptr), ls.mask.bytes),
		      "flsgpu_scan_filtered_to_host");
	}
	ls.cur_vector = 0;
}

// TableFilterSet -> device filters: the filter's column_index addresses column_ids (multi_file_function.hpp:356-379)
std::vector<flsgpu_filter> FastLanesReader::PhysicalFilters() const {
	std::vector<flsgpu_filter> out;
	for (const TableFilter& f : filters) {
		if (f.column_index >= column_ids.size() || IsFileRowNumberColumn(column_ids[f.column_index]) ||
		    column_ids[f.column_index] >= info.n_columns) {
			throw ReaderError(FLSGPU_ERR_INVALID, "Filter refers to a column that is not a projected physical column");
		}
		flsgpu_filter p;
		p.col = (uint32_t)column_ids[f.column_index];
		p.op = f.op;
		p.constant = f.constant;
		p.str_data = reinterpret_cast<const uint8_t*>(f.string_constant.data()); // owned by this->filters, alive for the scan
		p.str_len = f.string_constant.size();
		out.push_back(p);
	}
	return out;
}

// RowGroupFilter::BuildRowGroupsToScan / RowGroupMaySatisfyFilters (row_group_filter.cpp:80-199)
void FastLanesReader::BuildRowGroupFilter() {
	rowgroups_to_scan.assign(info.n_rowgroups, 1);
	const std::vector<flsgpu_filter> pf = PhysicalFilters();
	if (!pf.empty() && info.n_rowgroups) {
		Check(flsgpu_rowgroup_prune(table, pf.data(), (uint32_t)pf.size(), 0, info.n_rowgroups, rowgroups_to_scan.data()), "flsgpu_rowgroup_prune");
	}
	rowgroup_filter_built = true;
}

// FilterExecutor::Apply (filter_executor.cpp:19-60): the device already evaluated the filters for the whole batch; turn the
// mask bits of this chunk's rows into the selection vector the chunk is sliced with.
void FastLanesReader::ApplyFilters(FastLanesReadLocalState& ls, DataChunk& chunk, idx_t start_vector, idx_t count) const {
	chunk.sel.clear();
	if (!ls.has_mask) {
		return;
	}
	const idx_t rg_in_batch = ls.cur_rowgroup - ls.batch_begin;
	const uint32_t* words = static_cast<const uint32_t*>(ls.mask.ptr) + (rg_in_batch * info.vec_per_rowgroup + start_vector) * 32;
	for (idx_t w = 0; w < (count + 31) / 32; ++w) {
		uint32_t bits = words[w];
		while (bits) {
			const uint32_t i = (uint32_t)w * 32 + (uint32_t)__builtin_ctz(bits);
			bits &= bits - 1;
			if (i < count) {
				chunk.sel.push_back(i);
			}
		}
	}
	if (chunk.sel.size() == count) {
		chunk.sel.clear(); // nothing filtered: no slice (filter_executor.cpp:57)
	} else {
		chunk.SetCardinality(chunk.sel.size());
		if (chunk.sel.empty()) {
			chunk.sel.push_back(0); // keep "sliced" distinguishable from "identity"; count is 0
		}
	}
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
	if (batch.count > FLS_VEC_SZ && local_state.cur_vector != ~0ull &&
	    !VectorsJoin(local_state, (local_state.cur_rowgroup - local_state.batch_begin) * info.vec_per_rowgroup + local_state.cur_vector)) {
		batch.count = FLS_VEC_SZ;
	}
	batch.input_vector_count = (batch.count + FLS_VEC_SZ - 1) / FLS_VEC_SZ;
	return batch;
}

// FastLanesReader::DecodePhysicalColumns (fls_reader.cpp:436-472): the values are already decoded; point the chunk's
// vectors at rows [start_vector*1024, +count) of the current row group inside the staging buffers.
void FastLanesReader::DecodePhysicalColumns(FastLanesReadLocalState& ls, DataChunk& chunk, idx_t start_vector, idx_t count) {
	const idx_t rg_in_batch = ls.cur_rowgroup - ls.batch_begin;
	const idx_t first_elem = (rg_in_batch * info.vec_per_rowgroup + start_vector) * FLS_VEC_SZ;
	for (size_t i = 0; i < column_ids.size(); ++i) {
		const int64_t ri = ls.projection.output_to_reader_idx[i];
		if (ri < 0) {
			continue;
		}
		const ColumnDefinition& def = columns[ls.projection.reader_column_ids[ri]];
		Vector& v = chunk.data[i];
		v = Vector();
		v.type = def.type;
		if (ls.dict_form[ri]) { // Vector::Dictionary(dictionary of this row group, selection vector of the chunk's rows)
			v.vector_type = VectorType::DICTIONARY_VECTOR;
			v.sel = static_cast<const uint32_t*>(ls.staging[ri].ptr) + first_elem;
			v.data = static_cast<const uint8_t*>(ls.keys[ri].ptr) + ls.key_offsets[ri][rg_in_batch] * 16;
			v.dict_size = ls.key_offsets[ri][rg_in_batch + 1] - ls.key_offsets[ri][rg_in_batch];
		} else if (!ls.forms[ri].empty()) {
			FormedVector(ls, (size_t)ri, first_elem / FLS_VEC_SZ, count, v);
		} else {
			v.data = static_cast<const uint8_t*>(ls.staging[ri].ptr) + first_elem * def.type.physical_bytes;
		}
	}
}

// Do the two vectors of a would-be chunk continue each other in every compact column?  FLAT + FLAT side by side in the
// packed buffer, or CONSTANT / SEQUENCE halves where the second takes up where the first ends.  If not, the chunk is cut
// after its first vector (a chunk of 1024 rows, like the tail of a row group) so that every vector stays zero-copy.
bool FastLanesReader::VectorsJoin(const FastLanesReadLocalState& ls, idx_t first_vector) const {
	for (const auto& forms : ls.forms) {
		if (forms.empty()) {
			continue;
		}
		const flsgpu_vector_form &f0 = forms[first_vector], &f1 = forms[first_vector + 1];
		if (f0.form == FLS_VEC_FORM_FLAT || f1.form == FLS_VEC_FORM_FLAT) {
			if (f0.form != f1.form || f1.flat_slot != f0.flat_slot + 1) {
				return false;
			}
			continue;
		}
		const int64_t inc0 = f0.form == FLS_VEC_FORM_SEQUENCE ? f0.increment : 0, inc1 = f1.form == FLS_VEC_FORM_SEQUENCE ? f1.increment : 0;
		if (inc0 != inc1 || f1.start != f0.start + inc0 * (int64_t)FLS_VEC_SZ) {
			return false;
		}
	}
	return true;
}

// One chunk of a column scanned in compact form (its vectors join, see VectorsJoin): a zero-copy FLAT slice of the packed
// staging buffer, or a CONSTANT / SEQUENCE vector (constant_kernel.hpp:19-27 does the former for constant chunks).
void FastLanesReader::FormedVector(FastLanesReadLocalState& ls, size_t ri, idx_t first_vector, idx_t, Vector& v) const {
	const flsgpu_vector_form& f0 = ls.forms[ri][first_vector];
	if (f0.form == FLS_VEC_FORM_FLAT) {
		v.data = static_cast<const uint8_t*>(ls.staging[ri].ptr) + (size_t)f0.flat_slot * FLS_VEC_SZ * v.type.physical_bytes;
		return;
	}
	v.vector_type = f0.form == FLS_VEC_FORM_SEQUENCE ? VectorType::SEQUENCE_VECTOR : VectorType::CONSTANT_VECTOR;
	v.seq_start = f0.start;
	v.seq_increment = f0.form == FLS_VEC_FORM_SEQUENCE ? f0.increment : 0;
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
		chunk.data[i] = Vector();
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
	chunk.sel.clear();
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
		DecodePhysicalColumns(ls, chunk, ls.cur_vector, batch.count);
		const idx_t row_start = ls.row_group_base + batch.start_tuple;
		PopulateVirtualColumns(ls, chunk, row_start, batch.count);
		chunk.SetCardinality(batch.count);
		ApplyFilters(ls, chunk, ls.cur_vector, batch.count);
		ls.cur_vector += batch.input_vector_count;
		vectors_read += batch.input_vector_count;
		if (chunk.count == 0) {
			continue; // every row of the chunk was filtered out: serve the next one
		}
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
