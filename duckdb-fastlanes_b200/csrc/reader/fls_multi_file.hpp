// fls_multi_file.hpp — the multi-file side of read_fls: what the reference gets from DuckDB's MultiFileReader framework
// through FastLanesMultiFileInfo (reference src/reader/fls_multi_file_info.cpp:8-171, src/read_fls.cpp:32-46):
// globs / lists of files, named parameters, bind (first file's schema or union_by_name), cardinality estimate, MaxThreads,
// hand-out of files to workers, the `filename` and `file_row_number` virtual columns, missing columns of a union scan.
// DuckDB's types are replaced by the stand-ins of fls_reader.hpp; INTEGRATION.md shows the DuckDB-typed glue.
#pragma once
#include <map>

#include "fls_reader.hpp"

namespace flsreader {

constexpr idx_t COLUMN_IDENTIFIER_FILENAME = ~0ull - 2; // MultiFileReader::COLUMN_IDENTIFIER_FILENAME stand-in

// MultiFileOptions subset (duckdb/common/multi_file/multi_file_options.hpp) + the reader's own named parameters
struct MultiFileOptions {
	bool union_by_name = false;
	bool filename = false; // adds a `filename` VARCHAR column
};

struct BinderException : ReaderError {
	explicit BinderException(const std::string& msg) : ReaderError(FLSGPU_ERR_INVALID, msg) {
	}
};

// One column of the bound (global) schema and, per file, where it comes from.
struct BoundColumn {
	std::string name;
	LogicalType type;
};

// Per-thread scan state of a multi-file scan: the file being read + the per-file local state.
struct MultiFileLocalState {
	idx_t file_index = ~0ull;
	FastLanesReadLocalState reader_state;
	std::vector<int64_t> local_to_global; // projected output column -> position in the file reader's column_ids, -1: NULL column
	string_t filename_value {};
	std::vector<string_t> filename_rows;
};

struct MultiFileGlobalState {
	std::mutex lock; // MultiFileGlobalState::lock (multi_file_function.hpp:399-416)
	idx_t file_index = 0;
	std::vector<std::unique_ptr<FastLanesReadGlobalState>> reader_states; // one per file, created lazily
};

class FastLanesMultiFileInfo {
public:
	// read_fls(VARCHAR | LIST(VARCHAR)) (multi_file_reader.cpp:510-517): every pattern is glob-expanded, results sorted per
	// pattern like DuckDB's FileSystem::Glob; a pattern without a match -> IOException ("No files found that match")
	static std::vector<std::string> ExpandFiles(const std::vector<std::string>& patterns);

	// ParseOption (fls_multi_file_info.cpp:26-46): true when the key was consumed.  NULL values -> BinderException.
	bool ParseOption(const std::string& original_key, const std::string* value);

	// BindReader (fls_multi_file_info.cpp:70-83): opens the files, binds the global schema.
	void Bind(const std::vector<std::string>& files);

	const std::vector<BoundColumn>& Columns() const {
		return columns;
	}
	idx_t FileCount() const {
		return readers.size();
	}
	FastLanesReader& Reader(idx_t i) {
		return *readers[i];
	}
	const std::string& FileName(idx_t i) const {
		return files[i];
	}
	// GetCardinality (fls_multi_file_info.cpp:152-164)
	idx_t GetCardinality() const;
	// MaxThreads (fls_multi_file_info.cpp:99-110): ~0 = no limit (multiple files)
	idx_t MaxThreads() const;

	// projection of the scan: positions in Columns() (or COLUMN_IDENTIFIER_FILE_ROW_NUMBER / _FILENAME)
	std::vector<idx_t> column_ids;
	std::vector<TableFilter> filters; // column_index addresses column_ids

	// MultiFileScan (multi_file_function.hpp:591-616): fills `chunk` with the next non-empty chunk; false when the scan is done.
	bool Scan(MultiFileGlobalState& gstate, MultiFileLocalState& lstate, DataChunk& chunk);

	FastLanesFileReaderOptions options;
	MultiFileOptions file_options;

private:
	bool TryInitializeNextBatch(MultiFileGlobalState& gstate, MultiFileLocalState& lstate);
	void PrepareReader(idx_t file_index); // maps the global projection / filters onto the file's own columns
	std::vector<std::string> files;
	std::vector<std::unique_ptr<FastLanesReader>> readers;
	std::vector<BoundColumn> columns;
	std::vector<std::vector<int64_t>> global_to_file; // [file][global column] -> column of that file, -1 if absent
	std::vector<std::vector<int64_t>> file_local_to_global; // [file][projected output column] -> index into the reader's column_ids
	std::vector<uint8_t> prepared;
};

} // namespace flsreader
