// fls_reader_capi.cpp — see fls_reader_capi.h.
#include "fls_reader_capi.h"

#include <cstdio>
#include <cstring>
#include <thread>

#include "fls_multi_file.hpp"
#include "fls_reader.hpp"

using namespace flsreader;

struct flsreader_handle {
	std::unique_ptr<FastLanesReader> reader;
};

static thread_local std::string g_err;
const char* flsreader_last_error(void) {
	return g_err.c_str();
}

#define FLSR_TRY(...)                                                                                                  \
	try {                                                                                                              \
		__VA_ARGS__;                                                                                                     \
		return 0;                                                                                                      \
	} catch (const ReaderError& e) {                                                                                   \
		g_err = e.what();                                                                                              \
		return e.status;                                                                                               \
	} catch (const std::exception& e) {                                                                                \
		g_err = e.what();                                                                                              \
		return FLSGPU_ERR_INVALID;                                                                                     \
	}

int flsreader_open(const char* path, int file_row_number, uint64_t explicit_cardinality, uint64_t rowgroups_per_batch,
                   flsreader_handle** out) {
	FLSR_TRY({
		FastLanesFileReaderOptions o;
		o.file_row_number = file_row_number != 0;
		o.has_explicit_cardinality = explicit_cardinality != 0;
		o.explicit_cardinality = explicit_cardinality;
		if (rowgroups_per_batch) {
			o.rowgroups_per_batch = rowgroups_per_batch;
		}
		auto h = std::make_unique<flsreader_handle>();
		h->reader = std::make_unique<FastLanesReader>(path, o);
		*out = h.release();
	})
}
void flsreader_close(flsreader_handle* h) {
	delete h;
}
int flsreader_n_columns(const flsreader_handle* h) {
	return (int)h->reader->GetColumns().size();
}
int flsreader_column(const flsreader_handle* h, int i, char* name64, int* type_id, int* physical_bytes, int* width, int* scale) {
	const auto& cols = h->reader->GetColumns();
	if (i < 0 || (size_t)i >= cols.size()) {
		g_err = "column index out of range";
		return FLSGPU_ERR_INVALID;
	}
	std::snprintf(name64, 64, "%s", cols[i].name.c_str());
	*type_id = (int)cols[i].type.id;
	*physical_bytes = cols[i].type.physical_bytes;
	*width = cols[i].type.width;
	*scale = cols[i].type.scale;
	return 0;
}
uint64_t flsreader_cardinality(const flsreader_handle* h) {
	return h->reader->GetCardinality();
}
uint64_t flsreader_max_threads(const flsreader_handle* h) {
	return h->reader->MaxThreads();
}
uint64_t flsreader_total_tuples(const flsreader_handle* h) {
	return h->reader->GetTotalTuples();
}
double flsreader_progress(const flsreader_handle* h) {
	return h->reader->GetProgressInFile();
}
int flsreader_partition_stats(const flsreader_handle* h, uint64_t* row_start, uint64_t* count, uint64_t capacity, uint64_t* n) {
	std::vector<PartitionStatistics> ps;
	h->reader->GetPartitionStats(ps);
	*n = ps.size();
	for (size_t i = 0; i < ps.size() && i < capacity; ++i) {
		row_start[i] = ps[i].row_start;
		count[i] = ps[i].count;
	}
	return 0;
}
int flsreader_statistics(const flsreader_handle* h, const char* name, int* has_stats, int64_t* min_i, int64_t* max_i, double* min_d,
                         double* max_d) {
	FLSR_TRY({
		NumericStats s;
		*has_stats = h->reader->GetStatistics(name, s) ? 1 : 0;
		*min_i = s.min_i;
		*max_i = s.max_i;
		*min_d = s.min_d;
		*max_d = s.max_d;
	})
}

namespace {

// SELECT column_ids FROM read_fls(path) [WHERE filters]: drives TryInitializeScan / Scan on n_threads workers.
// Without filters rows land at their file row number; with filters the passing rows are appended in arrival order and
// their file row numbers go to row_ids_out.
void scan_impl(flsreader_handle* h, const uint64_t* column_ids, uint32_t n_cols, const flsreader_filter* filters, uint32_t n_filters,
               int n_threads, void* const* out_cols, int64_t* row_ids_out, flsreader_scan_stats* stats) {
	FastLanesReader& reader = *h->reader;
	// the driver appends file_row_number to the projection to know where each chunk belongs
	reader.column_ids.assign(column_ids, column_ids + n_cols);
	const bool place_rows = out_cols != nullptr || row_ids_out != nullptr; // a consumer that copies nothing needs no row numbers
	if (place_rows) {
		reader.column_ids.push_back(COLUMN_IDENTIFIER_FILE_ROW_NUMBER);
	}
	reader.filters.clear();
	reader.ResetScan();
	for (uint32_t i = 0; i < n_filters; ++i) {
		TableFilter f;
		f.column_index = filters[i].column_index;
		f.op = (flsgpu_cmp)filters[i].op;
		f.constant = filters[i].constant;
		if (filters[i].str_data && filters[i].str_len) {
			f.string_constant.assign(reinterpret_cast<const char*>(filters[i].str_data), filters[i].str_len);
		}
		reader.filters.push_back(f);
	}
	const auto& cols = reader.GetColumns();
	std::vector<uint32_t> widths(n_cols);
	std::vector<uint8_t> is_str(n_cols);
	for (uint32_t c = 0; c < n_cols; ++c) {
		if (column_ids[c] >= cols.size()) {
			throw ReaderError(FLSGPU_ERR_INVALID, "Projected column id out of range");
		}
		widths[c] = cols[column_ids[c]].type.physical_bytes;
		is_str[c] = cols[column_ids[c]].type.id == LogicalTypeId::VARCHAR || cols[column_ids[c]].type.id == LogicalTypeId::BLOB;
	}
	FastLanesReadGlobalState gstate;
	std::atomic<uint64_t> rows {0}, chunks {0}, maxrows {0}, dict_vectors {0}, constant_vectors {0}, sequence_vectors {0}, cursor {0};
	std::mutex err_mu;
	std::string err;
	int err_status = 0;
	const int nt = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)std::max(1, n_threads), reader.MaxThreads()));
	auto worker = [&]() {
		try {
			FastLanesReadLocalState lstate;
			DataChunk chunk;
			struct Tally { // folded into the shared counters once, when the worker is done
				uint64_t rows = 0, chunks = 0, maxrows = 0, dict = 0, constant = 0, sequence = 0;
			} tally;
			struct Fold {
				Tally& t;
				std::atomic<uint64_t>&rows, &chunks, &maxrows, &dict, &constant, &sequence;
				~Fold() {
					rows += t.rows;
					chunks += t.chunks;
					dict += t.dict;
					constant += t.constant;
					sequence += t.sequence;
					uint64_t m = maxrows.load();
					while (t.maxrows > m && !maxrows.compare_exchange_weak(m, t.maxrows)) {
					}
				}
			} fold {tally, rows, chunks, maxrows, dict_vectors, constant_vectors, sequence_vectors};
			for (;;) {
				{
					std::lock_guard<std::mutex> g(gstate.lock);
					if (!reader.TryInitializeScan(gstate, lstate)) {
						return;
					}
				}
				for (;;) {
					reader.Scan(gstate, lstate, chunk);
					if (chunk.count == 0) {
						break;
					}
					const int64_t* frn = place_rows ? static_cast<const int64_t*>(chunk.data[n_cols].data) : nullptr;
					const uint64_t base = row_ids_out ? cursor.fetch_add(chunk.count) : frn ? (uint64_t)frn[0] : 0;
					for (uint32_t c = 0; c < n_cols; ++c) {
						const Vector& v = chunk.data[c];
						tally.dict += v.vector_type == VectorType::DICTIONARY_VECTOR ? 1 : 0;
						tally.constant += v.vector_type == VectorType::CONSTANT_VECTOR ? 1 : 0;
						tally.sequence += v.vector_type == VectorType::SEQUENCE_VECTOR ? 1 : 0;
						if (!out_cols || !out_cols[c]) {
							continue; // the consumer takes the vector as it is (what DuckDB's operators do): nothing is copied
						}
						if (is_str[c]) {
							const flsreader_string_sink* sink = static_cast<const flsreader_string_sink*>(out_cols[c]);
							for (idx_t i = 0; i < chunk.count; ++i) {
								const string_t* sv = static_cast<const string_t*>(v.Row(chunk.RowIndex(i)));
								if (sv->GetSize() > sink->stride) {
									throw ReaderError(FLSGPU_ERR_INVALID, "string longer than the sink's stride");
								}
								std::memcpy(sink->bytes + (base + i) * sink->stride, sv->GetData(), sv->GetSize());
								sink->lengths[base + i] = sv->GetSize();
							}
						} else if (chunk.sel.empty() && v.vector_type == VectorType::FLAT_VECTOR) {
							std::memcpy(static_cast<uint8_t*>(out_cols[c]) + base * widths[c], v.data, chunk.count * widths[c]);
						} else {
							for (idx_t i = 0; i < chunk.count; ++i) {
								v.CopyRow(chunk.RowIndex(i), static_cast<uint8_t*>(out_cols[c]) + (base + i) * widths[c]);
							}
						}
					}
					for (idx_t i = 0; row_ids_out && i < chunk.count; ++i) {
						row_ids_out[base + i] = frn[chunk.RowIndex(i)];
					}
					tally.rows += chunk.count;
					tally.chunks += 1;
					tally.maxrows = std::max<uint64_t>(tally.maxrows, chunk.count);
				}
			}
		} catch (const ReaderError& e) {
			std::lock_guard<std::mutex> g(err_mu);
			err = e.what();
			err_status = e.status;
		} catch (const std::exception& e) {
			std::lock_guard<std::mutex> g(err_mu);
			err = e.what();
			err_status = FLSGPU_ERR_INVALID;
		}
	};
	std::vector<std::thread> th;
	for (int i = 0; i < nt; ++i) {
		th.emplace_back(worker);
	}
	for (auto& t : th) {
		t.join();
	}
	if (err_status) {
		throw ReaderError(err_status, err);
	}
	stats->rows = rows;
	stats->chunks = chunks;
	stats->max_chunk_rows = maxrows;
	stats->dictionary_vectors = dict_vectors;
	stats->constant_vectors = constant_vectors;
	stats->sequence_vectors = sequence_vectors;
	stats->assembled_vectors = 0; // chunks are split where the forms of their halves differ: nothing is put together on the host
	stats->pruned_rowgroups = reader.GetNPrunedRowGroups();
}

} // namespace

int flsreader_scan_all(flsreader_handle* h, const uint64_t* column_ids, uint32_t n_cols, int n_threads, void* const* out_cols,
                       uint64_t* rows_out, uint64_t* chunks_out, uint64_t* max_chunk_rows) {
	FLSR_TRY({
		flsreader_scan_stats st {};
		scan_impl(h, column_ids, n_cols, nullptr, 0, n_threads, out_cols, nullptr, &st);
		*rows_out = st.rows;
		*chunks_out = st.chunks;
		*max_chunk_rows = st.max_chunk_rows;
	})
}

int flsreader_evict(flsreader_handle* h) {
	FLSR_TRY({
		for (flsgpu_table* t : h->reader->DeviceTables()) {
			if (flsgpu_table_evict(t) != FLSGPU_OK) {
				throw ReaderError(FLSGPU_ERR_CUDA, flsgpu_last_error());
			}
		}
	})
}

int flsreader_set_option(flsreader_handle* h, const char* key, uint64_t value) {
	FLSR_TRY({
		const std::string k = key ? key : "";
		if (k == "compressed_vectors") {
			h->reader->SetCompressedVectors(value != 0);
		} else {
			throw ReaderError(FLSGPU_ERR_INVALID, "unknown reader option: " + k);
		}
	})
}

int flsreader_scan(flsreader_handle* h, const uint64_t* column_ids, uint32_t n_cols, const flsreader_filter* filters, uint32_t n_filters,
                   int n_threads, void* const* out_cols, int64_t* row_ids_out, flsreader_scan_stats* stats) {
	FLSR_TRY({ scan_impl(h, column_ids, n_cols, filters, n_filters, n_threads, out_cols, n_filters ? row_ids_out : nullptr, stats); })
}

// ---- multi-file ---------------------------------------------------------------------------------------------------------
struct flsmulti_handle {
	FastLanesMultiFileInfo info;
};

int flsmulti_open(const char* const* patterns, uint32_t n_patterns, const char* const* option_keys, const char* const* option_values,
                  uint32_t n_options, flsmulti_handle** out) {
	FLSR_TRY({
		auto h = std::make_unique<flsmulti_handle>();
		for (uint32_t i = 0; i < n_options; ++i) {
			std::string v;
			if (option_values[i]) {
				v = option_values[i];
			}
			if (!h->info.ParseOption(option_keys[i], option_values[i] ? &v : nullptr)) {
				throw BinderException(std::string("Invalid named parameter \"") + option_keys[i] + "\" for function read_fls");
			}
		}
		std::vector<std::string> pats(patterns, patterns + n_patterns);
		h->info.Bind(FastLanesMultiFileInfo::ExpandFiles(pats));
		*out = h.release();
	})
}
void flsmulti_close(flsmulti_handle* h) {
	delete h;
}
int flsmulti_n_files(const flsmulti_handle* h) {
	return (int)h->info.FileCount();
}
int flsmulti_file(const flsmulti_handle* h, int i, char* name256, uint64_t* n_tuples) {
	if (i < 0 || (idx_t)i >= h->info.FileCount()) {
		g_err = "file index out of range";
		return FLSGPU_ERR_INVALID;
	}
	std::snprintf(name256, 256, "%s", h->info.FileName(i).c_str());
	*n_tuples = const_cast<flsmulti_handle*>(h)->info.Reader(i).GetTotalTuples();
	return 0;
}
int flsmulti_n_columns(const flsmulti_handle* h) {
	return (int)h->info.Columns().size();
}
int flsmulti_column(const flsmulti_handle* h, int i, char* name64, int* type_id, int* physical_bytes, int* width, int* scale) {
	const auto& cols = h->info.Columns();
	if (i < 0 || (size_t)i >= cols.size()) {
		g_err = "column index out of range";
		return FLSGPU_ERR_INVALID;
	}
	std::snprintf(name64, 64, "%s", cols[i].name.c_str());
	*type_id = (int)cols[i].type.id;
	*physical_bytes = cols[i].type.physical_bytes;
	*width = cols[i].type.width;
	*scale = cols[i].type.scale;
	return 0;
}
uint64_t flsmulti_cardinality(const flsmulti_handle* h) {
	return h->info.GetCardinality();
}
uint64_t flsmulti_max_threads(const flsmulti_handle* h) {
	return h->info.MaxThreads();
}

int flsmulti_scan(flsmulti_handle* h, const uint64_t* column_ids, uint32_t n_cols, const flsreader_filter* filters, uint32_t n_filters,
                  int n_threads, void* const* out_cols, uint8_t* const* valid_out, uint32_t* file_ids_out, int64_t* row_ids_out,
                  flsreader_scan_stats* stats) {
	FLSR_TRY({
		FastLanesMultiFileInfo& info = h->info;
		// a fresh handle per scan keeps this simple: the projection is bound once
		info.column_ids.assign(column_ids, column_ids + n_cols);
		info.column_ids.push_back(COLUMN_IDENTIFIER_FILE_ROW_NUMBER);
		info.filters.clear();
		for (uint32_t i = 0; i < n_filters; ++i) {
			TableFilter f;
			f.column_index = filters[i].column_index;
			f.op = (flsgpu_cmp)filters[i].op;
			f.constant = filters[i].constant;
			if (filters[i].str_data && filters[i].str_len) {
				f.string_constant.assign(reinterpret_cast<const char*>(filters[i].str_data), filters[i].str_len);
			}
			info.filters.push_back(f);
		}
		const auto& cols = info.Columns();
		std::vector<uint32_t> widths(n_cols);
		std::vector<uint8_t> is_str(n_cols);
		for (uint32_t c = 0; c < n_cols; ++c) {
			if (column_ids[c] >= cols.size()) {
				throw ReaderError(FLSGPU_ERR_INVALID, "Projected column id out of range");
			}
			widths[c] = cols[column_ids[c]].type.physical_bytes;
			is_str[c] = cols[column_ids[c]].type.id == LogicalTypeId::VARCHAR || cols[column_ids[c]].type.id == LogicalTypeId::BLOB;
		}
		MultiFileGlobalState gstate;
		std::atomic<uint64_t> rows {0}, chunks {0}, maxrows {0}, dict_vectors {0}, cursor {0};
		std::mutex err_mu;
		std::string err;
		int err_status = 0;
		const uint64_t mt = info.MaxThreads();
		const int nt = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)std::max(1, n_threads), mt));
		auto worker = [&]() {
			try {
				MultiFileLocalState lstate;
				DataChunk chunk;
				while (info.Scan(gstate, lstate, chunk)) {
					const int64_t* frn = static_cast<const int64_t*>(chunk.data[n_cols].data);
					const uint64_t base = cursor.fetch_add(chunk.count);
					for (uint32_t c = 0; c < n_cols; ++c) {
						const Vector& v = chunk.data[c];
						dict_vectors += v.vector_type == VectorType::DICTIONARY_VECTOR ? 1 : 0;
						for (idx_t i = 0; i < chunk.count; ++i) {
							if (valid_out && valid_out[c]) {
								valid_out[c][base + i] = v.all_null ? 0 : 1;
							}
							if (v.all_null) {
								continue;
							}
							if (is_str[c]) {
								const flsreader_string_sink* sink = static_cast<const flsreader_string_sink*>(out_cols[c]);
								const string_t* sv = static_cast<const string_t*>(v.Row(chunk.RowIndex(i)));
								if (sv->GetSize() > sink->stride) {
									throw ReaderError(FLSGPU_ERR_INVALID, "string longer than the sink's stride");
								}
								std::memcpy(sink->bytes + (base + i) * sink->stride, sv->GetData(), sv->GetSize());
								sink->lengths[base + i] = sv->GetSize();
							} else {
								v.CopyRow(chunk.RowIndex(i), static_cast<uint8_t*>(out_cols[c]) + (base + i) * widths[c]);
							}
						}
					}
					for (idx_t i = 0; i < chunk.count; ++i) {
						row_ids_out[base + i] = frn[chunk.RowIndex(i)];
						file_ids_out[base + i] = (uint32_t)lstate.file_index;
					}
					rows += chunk.count;
					chunks += 1;
					uint64_t m = maxrows.load();
					while (chunk.count > m && !maxrows.compare_exchange_weak(m, chunk.count)) {
					}
				}
			} catch (const ReaderError& e) {
				std::lock_guard<std::mutex> g(err_mu);
				err = e.what();
				err_status = e.status;
			} catch (const std::exception& e) {
				std::lock_guard<std::mutex> g(err_mu);
				err = e.what();
				err_status = FLSGPU_ERR_INVALID;
			}
		};
		std::vector<std::thread> th;
		for (int i = 0; i < nt; ++i) {
			th.emplace_back(worker);
		}
		for (auto& t : th) {
			t.join();
		}
		if (err_status) {
			throw ReaderError(err_status, err);
		}
		stats->rows = rows;
		stats->chunks = chunks;
		stats->max_chunk_rows = maxrows;
		stats->dictionary_vectors = dict_vectors;
		stats->pruned_rowgroups = 0;
		for (idx_t f = 0; f < info.FileCount(); ++f) {
			stats->pruned_rowgroups += info.Reader(f).GetNPrunedRowGroups();
		}
	})
}
