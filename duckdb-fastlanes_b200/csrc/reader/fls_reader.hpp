// fls_reader.hpp — host-side scan driver above the C ABI: the read_fls reader of the extension, re-written for a
// batched GPU decode.  Mirrors the reference's FastLanesReader (src/include/reader/fls_reader.hpp:27-104,
// src/reader/fls_reader.cpp) method for method — Initialize, TryInitializeScan, TryAssignNextRowGroup,
// BuildReaderProjectionPlan, GetScanBatch, DecodePhysicalColumns, PopulateVirtualColumns, Scan, GetProgressInFile,
// GetPartitionStats — with DuckDB's types replaced by the minimal stand-ins below so that it builds and is tested
// without a DuckDB tree.  The DuckDB-typed glue on top (Vector / DataChunk plumbing) is shown in INTEGRATION.md.
//
// What changes against the reference: a worker is handed a BATCH of row groups (not one), the batch is staged,
// decoded and copied back by ONE pipelined call (flsgpu_scan_columns_to_host), and Scan() serves 2048-row chunks
// as zero-copy slices of the pinned staging buffers (FlatVector::SetData style) instead of decoding two vectors per call.
#pragma once
#include <atomic>
#include <cstdint>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "flsgpu.h"

namespace flsreader {

using idx_t = uint64_t;
constexpr idx_t STANDARD_VECTOR_SIZE = 2048; // duckdb/src/include/duckdb/common/vector_size.hpp:16
constexpr idx_t COLUMN_IDENTIFIER_FILE_ROW_NUMBER = ~0ull - 1; // virtual column (fls_multi_file_info.cpp:166-171)

// DuckDB LogicalTypeId subset the reader binds (src/reader/translation_utils.cpp:5-48)
enum class LogicalTypeId : uint8_t { INVALID, BOOLEAN, TINYINT, SMALLINT, INTEGER, BIGINT, UBIGINT, FLOAT, DOUBLE, DATE, TIMESTAMP, DECIMAL, VARCHAR, BLOB };

// duckdb::string_t (duckdb/src/include/duckdb/common/types/string_type.hpp:22-60,236-247): what the device writes for
// VARCHAR / BLOB values; non-inlined strings point into the reader's pinned string heap.
struct string_t {
	static constexpr uint32_t INLINE_LENGTH = 12;
	uint32_t length;
	char body[12]; // inlined[12] | prefix[4] + pointer
	uint32_t GetSize() const {
		return length;
	}
	const char* GetData() const {
		if (length <= INLINE_LENGTH) {
			return body;
		}
		const char* p;
		__builtin_memcpy(&p, body + 4, sizeof(p));
		return p;
	}
};
static_assert(sizeof(string_t) == 16, "string_t must be 16 bytes");

struct LogicalType {
	LogicalTypeId id = LogicalTypeId::INVALID;
	uint8_t width = 0, scale = 0; // DECIMAL
	uint8_t physical_bytes = 0;   // size of one value in a chunk vector
};

struct ColumnDefinition { // MultiFileColumnDefinition
	std::string name;
	LogicalType type;
};

// FastLanesFileReaderOptions (fls_multi_file_info.hpp:16-23)
struct FastLanesFileReaderOptions {
	bool has_explicit_cardinality = false;
	idx_t explicit_cardinality = 0;
	bool file_row_number = false;
	// GPU-side knobs (no reference counterpart)
	idx_t rowgroups_per_batch = 32;
	// hand vectors the file marks CONSTANT / arithmetic SEQUENCE on in that form (the reference does it for constant chunks:
	// constant_kernel.hpp:19-27) instead of decoding and copying them; the extension's setting fastlanes_compressed_vectors
	bool compressed_vectors = true;
};

// Stand-in for duckdb::Vector / DataChunk; data is owned by the reader's staging buffers (FlatVector::SetData /
// Vector::Dictionary style hand-off, dictionary_kernel.hpp:80-101).
enum class VectorType : uint8_t { FLAT_VECTOR, DICTIONARY_VECTOR, CONSTANT_VECTOR, SEQUENCE_VECTOR };
struct Vector {
	LogicalType type;
	VectorType vector_type = VectorType::FLAT_VECTOR;
	const void* data = nullptr;    // FLAT: the chunk's values; DICTIONARY: the dictionary's values (dict_size of them)
	const uint32_t* sel = nullptr; // DICTIONARY: selection vector, one entry per row of the (unsliced) chunk
	idx_t dict_size = 0;
	bool all_null = false;         // CONSTANT NULL vector: a union_by_name column this file does not have
	int64_t seq_start = 0, seq_increment = 0; // CONSTANT: the value (physical word); SEQUENCE: start + i * increment
	// value of (unsliced) row i as a pointer to its physical representation (FLAT / DICTIONARY vectors)
	const void* Row(idx_t i) const {
		const idx_t k = vector_type == VectorType::DICTIONARY_VECTOR ? sel[i] : i;
		return static_cast<const uint8_t*>(data) + k * type.physical_bytes;
	}
	// value of (unsliced) row i copied to dst, whatever the vector type (numeric columns)
	void CopyRow(idx_t i, void* dst) const {
		if (vector_type == VectorType::CONSTANT_VECTOR || vector_type == VectorType::SEQUENCE_VECTOR) {
			const uint64_t v = (uint64_t)seq_start + (uint64_t)seq_increment * (vector_type == VectorType::SEQUENCE_VECTOR ? i : 0);
			__builtin_memcpy(dst, &v, type.physical_bytes); // little endian: the low bytes are the value
		} else {
			__builtin_memcpy(dst, Row(i), type.physical_bytes);
		}
	}
};
struct DataChunk {
	std::vector<Vector> data;
	idx_t count = 0;
	// DataChunk::Slice(sel, count) after FilterExecutor::Apply (filter_executor.cpp:57-59): row i of the chunk is row
	// sel[i] of the vectors; empty = identity
	std::vector<uint32_t> sel;
	void SetCardinality(idx_t n) {
		count = n;
	}
	idx_t RowIndex(idx_t i) const {
		return sel.empty() ? i : sel[i];
	}
};

// ConstantFilter entry of a TableFilterSet (BaseFileReader::filters, base_file_reader.hpp:36-54): keyed by the position
// of the column in column_ids; the constant is the physical value of the column's type (flsgpu_filter), or the bytes of
// a VARCHAR / BLOB constant (string_constant; compared like DuckDB's string_t, string_type.hpp:155-217).
struct TableFilter {
	idx_t column_index = 0; // index into column_ids
	flsgpu_cmp op = FLSGPU_CMP_EQ;
	uint64_t constant = 0;
	std::string string_constant; // filters on string columns only
};

// FastLanesReaderProjectionPlan (fls_multi_file_info.hpp:48-55)
struct FastLanesReaderProjectionPlan {
	std::vector<uint32_t> reader_column_ids;       // physical column ids passed to the decoder
	std::vector<int64_t> output_to_reader_idx;     // output column -> index into reader_column_ids, -1 for virtual columns
	bool has_physical_columns = false;
};

struct FastLanesReadGlobalState { // fls_multi_file_info.hpp:79-86 (+ the multi-file lock TryInitializeScan runs under)
	idx_t next_rowgroup = 0;
	idx_t next_batch = 0; // batches handed out so far: batch k goes to device k mod G
	std::mutex lock;
};

struct PinnedBuffer {
	void* ptr = nullptr;
	uint64_t bytes = 0;
	~PinnedBuffer();
	void Reserve(uint64_t n);
};

struct FastLanesReadLocalState { // fls_multi_file_info.hpp:57-77
	idx_t n_tuples = 0;       // tuples of the current row group
	idx_t cur_rowgroup = 0;   // row group being served
	idx_t cur_vector = 0;     // next vector of that row group
	idx_t row_group_base = 0; // absolute row offset of the current row group
	FastLanesReaderProjectionPlan projection;
	// batch decoded by the last TryInitializeScan
	idx_t batch_begin = 0, batch_end = 0; // row groups [begin, end)
	idx_t device_slot = 0;                // which of the reader's per-device table handles decodes this batch
	std::vector<PinnedBuffer> staging;    // one per reader column: batch_rowgroups * vec_per_rowgroup * 1024 values
	// string columns: string heap, dictionary keys (dictionary form) and the first key of every row group of the batch
	std::vector<PinnedBuffer> heaps, keys;
	std::vector<uint8_t> dict_form;
	// compact scans: the form of every vector of the batch per reader column (empty: scanned in place)
	std::vector<std::vector<flsgpu_vector_form>> forms;
	std::vector<std::vector<idx_t>> key_offsets;
	PinnedBuffer mask;                    // selection mask of the pushed-down filters for the batch
	bool has_mask = false;
	std::vector<int64_t> row_numbers;     // file_row_number chunk buffer
};

struct PartitionStatistics { // duckdb PartitionStatistics: exact row ranges per row group
	idx_t row_start = 0;
	idx_t count = 0;
};

struct NumericStats {
	bool has_stats = false;
	int64_t min_i = 0, max_i = 0; // integers (sign-extended physical words; unsigned: reinterpret)
	double min_d = 0, max_d = 0;  // FLOAT / DOUBLE
};

class FastLanesReader {
public:
	explicit FastLanesReader(const std::string& path, const FastLanesFileReaderOptions& options = {});
	~FastLanesReader();
	FastLanesReader(const FastLanesReader&) = delete;

	std::string GetReaderType() const {
		return "FastLanes";
	}
	// filled by the caller before the scan, like BaseFileReader::column_ids (base_file_reader.hpp:36-54)
	std::vector<idx_t> column_ids;
	std::vector<TableFilter> filters; // pushed-down constant comparisons (read_fls.cpp:39-42: filter_pushdown = true)
	const std::vector<ColumnDefinition>& GetColumns() const {
		return columns;
	}

	// a new scan over the same reader (new column_ids / filters): forget the row-group filter and the progress counter
	void ResetScan() {
		rowgroup_filter_built = false;
		rowgroups_to_scan.clear();
		vectors_read = 0;
	}
	bool TryInitializeScan(FastLanesReadGlobalState& gstate, FastLanesReadLocalState& lstate);
	void Scan(FastLanesReadGlobalState& gstate, FastLanesReadLocalState& lstate, DataChunk& chunk);
	double GetProgressInFile() const;
	bool GetStatistics(const std::string& name, NumericStats& out) const;
	void GetPartitionStats(std::vector<PartitionStatistics>& result) const;
	idx_t GetCardinality() const; // FastLanesMultiFileInfo::GetCardinality (fls_multi_file_info.cpp:152-164)
	idx_t MaxThreads() const;     // fls_multi_file_info.cpp:99-110: #row groups (here: #batches)

	idx_t GetNRowGroups() const;
	idx_t GetNTuples(idx_t row_group_idx) const;
	idx_t GetNVectors(idx_t row_group_idx) const;
	size_t GetTotalTuples() const {
		return total_tuples;
	}
	idx_t GetNPrunedRowGroups() const {
		idx_t n = 0;
		for (uint8_t k : rowgroups_to_scan) {
			n += k ? 0 : 1;
		}
		return n;
	}
	size_t GetTotalVectors() const {
		return total_vectors;
	}
	const FastLanesFileReaderOptions& GetOptions() const {
		return reader_options;
	}
	flsgpu_table* Table() const {
		return table;
	}
	const std::vector<flsgpu_table*>& DeviceTables() const {
		return device_tables;
	}
	void SetCompressedVectors(bool on) {
		reader_options.compressed_vectors = on;
	}

private:
	struct ScanBatch {
		idx_t start_tuple = 0;
		idx_t count = 0;
		idx_t input_vector_count = 0;
		bool IsValid() const {
			return input_vector_count > 0;
		}
	};
	void Initialize();
	bool TryAssignNextRowGroup(FastLanesReadGlobalState& global_state, FastLanesReadLocalState& local_state);
	void DecodeRowGroupBatch(FastLanesReadLocalState& local_state);
	ScanBatch GetScanBatch(const FastLanesReadLocalState& local_state) const;
	void DecodePhysicalColumns(FastLanesReadLocalState& local_state, DataChunk& chunk, idx_t start_vector, idx_t count);
	bool VectorsJoin(const FastLanesReadLocalState& local_state, idx_t first_vector) const;
	void FormedVector(FastLanesReadLocalState& local_state, size_t reader_column, idx_t first_vector, idx_t count, Vector& v) const;
	void PopulateVirtualColumns(FastLanesReadLocalState& local_state, DataChunk& chunk, idx_t row_start, idx_t count);
	void ApplyFilters(FastLanesReadLocalState& local_state, DataChunk& chunk, idx_t start_vector, idx_t count) const;
	std::vector<flsgpu_filter> PhysicalFilters() const;
	void BuildRowGroupFilter();
	FastLanesReaderProjectionPlan BuildReaderProjectionPlan() const;
	bool IsFileRowNumberColumn(idx_t column_id) const;

	flsgpu_table* table = nullptr;
	// Single-process multi-GPU (the reference runs one worker per row group inside one process,
	// fls_multi_file_info.cpp:99-110): one handle per device of flsgpu_init's list over the same host image
	// (device_tables[0] == table).  Row groups are independent, so a batch simply goes to the next device in turn.
	std::vector<flsgpu_table*> device_tables;
	flsgpu_table_info info {};
	std::vector<ColumnDefinition> columns;
	std::vector<int32_t> column_out_types;   // flsgpu_out_type per physical column
	std::vector<idx_t> row_group_offsets;    // first tuple of each row group
	std::vector<fls_rowgroup_desc> row_groups;
	std::vector<uint8_t> rowgroups_to_scan;  // RowGroupFilter (row_group_filter.cpp:80-94): 0 = pruned by its zone maps
	bool rowgroup_filter_built = false;
	size_t total_vectors = 0; // the reference leaves these uninitialised; zeroed here
	size_t total_tuples = 0;
	std::atomic<idx_t> vectors_read {0};
	FastLanesFileReaderOptions reader_options;
};

// TranslateUtils::TranslateType (src/reader/translation_utils.cpp:5-48): every unsigned width reads back as UBIGINT.
LogicalType TranslateType(const fls_column_desc& col, int32_t* out_type);

struct ReaderError : std::runtime_error {
	int status;
	ReaderError(int status_p, const std::string& msg) : std::runtime_error(msg), status(status_p) {
	}
};

} // namespace flsreader
