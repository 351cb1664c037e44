// fls_multi_file.cpp — see fls_multi_file.hpp.  Every method cites the reference / DuckDB code it stands in for.
#include "fls_multi_file.hpp"

#include <glob.h>

#include <algorithm>
#include <cstring>

namespace flsreader {

namespace {
std::string Lower(std::string s) {
	std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); });
	return s;
}
bool ParseBool(const std::string& v, const std::string& key) {
	const std::string l = Lower(v);
	if (l == "true" || l == "1" || l == "t") {
		return true;
	}
	if (l == "false" || l == "0" || l == "f") {
		return false;
	}
	throw BinderException("Invalid BOOLEAN value for " + key + ": " + v);
}
bool SameType(const LogicalType& a, const LogicalType& b) {
	return a.id == b.id && a.width == b.width && a.scale == b.scale && a.physical_bytes == b.physical_bytes;
}
string_t MakeString(const std::string& s) {
	string_t r {};
	r.length = (uint32_t)s.size();
	if (s.size() <= string_t::INLINE_LENGTH) {
		std::memcpy(r.body, s.data(), s.size());
	} else {
		std::memcpy(r.body, s.data(), 4);
		const char* p = s.data();
		std::memcpy(r.body + 4, &p, sizeof(p));
	}
	return r;
}
} // namespace

// MultiFileReader::CreateFileList + FileSystem::Glob: patterns with wildcards are expanded and sorted, plain paths pass.
std::vector<std::string> FastLanesMultiFileInfo::ExpandFiles(const std::vector<std::string>& patterns) {
	std::vector<std::string> out;
	for (const std::string& pat : patterns) {
		if (pat.find_first_of("*?[") == std::string::npos) {
			out.push_back(pat);
			continue;
		}
		glob_t g {};
		const int rc = glob(pat.c_str(), 0, nullptr, &g);
		std::vector<std::string> found;
		for (size_t i = 0; rc == 0 && i < g.gl_pathc; ++i) {
			found.emplace_back(g.gl_pathv[i]);
		}
		globfree(&g);
		if (found.empty()) { // duckdb: IOException "No files found that match the pattern"
			throw ReaderError(FLSGPU_ERR_IO, "No files found that match the pattern \"" + pat + "\"");
		}
		std::sort(found.begin(), found.end());
		out.insert(out.end(), found.begin(), found.end());
	}
	if (out.empty()) {
		throw ReaderError(FLSGPU_ERR_IO, "read_fls needs at least one file");
	}
	return out;
}

// FastLanesMultiFileInfo::ParseOption (fls_multi_file_info.cpp:26-46) + the MultiFileReader options the scan honours
bool FastLanesMultiFileInfo::ParseOption(const std::string& original_key, const std::string* value) {
	const std::string key = Lower(original_key);
	const bool known = key == "explicit_cardinality" || key == "file_row_number" || key == "union_by_name" || key == "filename";
	if (known && !value) {
		throw BinderException("Cannot use NULL as argument to " + original_key);
	}
	if (key == "explicit_cardinality") {
		size_t pos = 0;
		unsigned long long v = 0;
		try {
			v = std::stoull(*value, &pos);
		} catch (const std::exception&) {
			pos = 0;
		}
		if (pos != value->size() || value->empty() || (*value)[0] == '-') {
			throw BinderException("explicit_cardinality expects a UBIGINT, got " + *value);
		}
		options.explicit_cardinality = v;
		options.has_explicit_cardinality = v != 0;
		return true;
	}
	if (key == "file_row_number") {
		options.file_row_number = ParseBool(*value, original_key);
		return true;
	}
	if (key == "union_by_name") {
		file_options.union_by_name = ParseBool(*value, original_key);
		return true;
	}
	if (key == "filename") {
		file_options.filename = ParseBool(*value, original_key);
		return true;
	}
	return false;
}

// BindReader (fls_multi_file_info.cpp:70-83): MultiFileReader::BindReader takes the first file's schema,
// BindUnionReader the union of all files' columns by name (first appearance order).
void FastLanesMultiFileInfo::Bind(const std::vector<std::string>& files_p) {
	files = files_p;
	readers.clear();
	columns.clear();
	FastLanesFileReaderOptions per_file = options;
	per_file.file_row_number = false; // the virtual column is bound once, on the global schema
	for (const std::string& f : files) {
		readers.push_back(std::make_unique<FastLanesReader>(f, per_file));
	}
	const size_t n_schema_files = file_options.union_by_name ? readers.size() : 1;
	for (size_t fi = 0; fi < n_schema_files; ++fi) {
		for (const ColumnDefinition& c : readers[fi]->GetColumns()) {
			auto it = std::find_if(columns.begin(), columns.end(), [&](const BoundColumn& b) { return b.name == c.name; });
			if (it == columns.end()) {
				columns.push_back({c.name, c.type});
			} else if (!SameType(it->type, c.type)) {
				throw BinderException("union_by_name: column \"" + c.name + "\" has different types in different files");
			}
		}
	}
	global_to_file.assign(readers.size(), std::vector<int64_t>(columns.size(), -1));
	for (size_t fi = 0; fi < readers.size(); ++fi) {
		const auto& fc = readers[fi]->GetColumns();
		for (size_t g = 0; g < columns.size(); ++g) {
			for (size_t k = 0; k < fc.size(); ++k) {
				if (fc[k].name == columns[g].name) {
					if (!SameType(fc[k].type, columns[g].type)) {
						throw BinderException("column \"" + fc[k].name + "\" of " + files[fi] + " does not have the bound type");
					}
					global_to_file[fi][g] = (int64_t)k;
				}
			}
			if (global_to_file[fi][g] < 0 && !file_options.union_by_name) { // MultiFileColumnMapper: missing column without union_by_name
				throw ReaderError(FLSGPU_ERR_INVALID, "file " + files[fi] + " does not contain column \"" + columns[g].name + "\"");
			}
		}
	}
	if (file_options.filename) {
		BoundColumn b;
		b.name = "filename";
		b.type.id = LogicalTypeId::VARCHAR;
		b.type.physical_bytes = 16;
		columns.push_back(b);
	}
	if (options.file_row_number) { // EnsureFileRowNumberColumn / GetVirtualColumns (fls_multi_file_info.cpp:166-171)
		BoundColumn b;
		b.name = "file_row_number";
		b.type.id = LogicalTypeId::BIGINT;
		b.type.physical_bytes = 8;
		columns.push_back(b);
	}
	file_local_to_global.assign(readers.size(), {});
	prepared.assign(readers.size(), 0);
}

// FastLanesMultiFileInfo::GetCardinality (fls_multi_file_info.cpp:152-164)
idx_t FastLanesMultiFileInfo::GetCardinality() const {
	if (options.explicit_cardinality) {
		return options.explicit_cardinality;
	}
	const idx_t initial = readers.empty() ? 0 : readers[0]->GetTotalTuples();
	if (readers.size() == 1 && initial) {
		return initial;
	}
	return std::max<idx_t>(initial, 42) * readers.size(); // fallback when the first file does not say much
}

// FastLanesMultiFileInfo::MaxThreads (fls_multi_file_info.cpp:99-110)
idx_t FastLanesMultiFileInfo::MaxThreads() const {
	if (readers.size() > 1) {
		return ~0ull; // multiple files: no limit, so that a small first file does not serialise the scan
	}
	return readers.empty() ? 1 : std::max<idx_t>(readers[0]->MaxThreads(), 1);
}

// MultiFileColumnMapper::CreateMapping: the global projection and filters in terms of file `fi`'s own columns.
void FastLanesMultiFileInfo::PrepareReader(idx_t fi) {
	if (prepared[fi]) {
		return;
	}
	FastLanesReader& r = *readers[fi];
	r.column_ids.clear();
	r.filters.clear();
	r.ResetScan();
	std::vector<int64_t>& map = file_local_to_global[fi];
	map.assign(column_ids.size(), -1);
	const size_t n_phys = columns.size() - (file_options.filename ? 1 : 0) - (options.file_row_number ? 1 : 0);
	for (size_t i = 0; i < column_ids.size(); ++i) {
		const idx_t g = column_ids[i];
		const bool is_filename = g == COLUMN_IDENTIFIER_FILENAME || (file_options.filename && g == n_phys);
		const bool is_frn = g == COLUMN_IDENTIFIER_FILE_ROW_NUMBER ||
		                    (options.file_row_number && g == n_phys + (file_options.filename ? 1 : 0));
		if (is_filename) {
			map[i] = -2;
		} else if (is_frn) {
			map[i] = (int64_t)r.column_ids.size();
			r.column_ids.push_back(COLUMN_IDENTIFIER_FILE_ROW_NUMBER);
		} else if (g >= n_phys) {
			throw ReaderError(FLSGPU_ERR_INVALID, "Projected column id out of range");
		} else if (global_to_file[fi][g] >= 0) {
			map[i] = (int64_t)r.column_ids.size();
			r.column_ids.push_back((idx_t)global_to_file[fi][g]);
		} // else: the file lacks the column -> constant NULL
	}
	for (const TableFilter& f : filters) {
		if (f.column_index >= column_ids.size() || map[f.column_index] == -2) {
			throw ReaderError(FLSGPU_ERR_INVALID, "Filter refers to a column that is not a projected physical column");
		}
		if (map[f.column_index] < 0) {
			// comparison with a NULL column is never true: nothing of this file qualifies
			prepared[fi] = 2;
			return;
		}
		TableFilter t = f;
		t.column_index = (idx_t)map[f.column_index];
		r.filters.push_back(t);
	}
	prepared[fi] = 1;
}

// MultiFileFunction::TryInitializeNextBatch (multi_file_function.hpp:397-472): under the global lock, ask the current
// file's reader for work; when it has none left, move on to the next file.
bool FastLanesMultiFileInfo::TryInitializeNextBatch(MultiFileGlobalState& gstate, MultiFileLocalState& lstate) {
	std::lock_guard<std::mutex> g(gstate.lock);
	if (gstate.reader_states.size() < readers.size()) {
		gstate.reader_states.resize(readers.size());
	}
	while (gstate.file_index < readers.size()) {
		const idx_t fi = gstate.file_index;
		PrepareReader(fi);
		if (prepared[fi] == 1) {
			if (!gstate.reader_states[fi]) {
				gstate.reader_states[fi] = std::make_unique<FastLanesReadGlobalState>();
			}
			if (readers[fi]->TryInitializeScan(*gstate.reader_states[fi], lstate.reader_state)) {
				if (lstate.file_index != fi) {
					lstate.file_index = fi;
					lstate.local_to_global = file_local_to_global[fi];
					lstate.filename_value = MakeString(files[fi]);
					lstate.filename_rows.assign(STANDARD_VECTOR_SIZE, lstate.filename_value);
				}
				return true;
			}
		}
		gstate.file_index++;
	}
	return false;
}

bool FastLanesMultiFileInfo::Scan(MultiFileGlobalState& gstate, MultiFileLocalState& lstate, DataChunk& chunk) {
	DataChunk file_chunk;
	for (;;) {
		if (lstate.file_index != ~0ull) {
			FastLanesReader& r = *readers[lstate.file_index];
			r.Scan(*gstate.reader_states[lstate.file_index], lstate.reader_state, file_chunk);
			if (file_chunk.count > 0) {
				chunk.data.resize(column_ids.size());
				for (size_t i = 0; i < column_ids.size(); ++i) {
					const int64_t j = lstate.local_to_global[i];
					Vector& v = chunk.data[i];
					if (j >= 0) {
						v = file_chunk.data[j];
					} else if (j == -2) { // filename: the same string for every row of the file
						v = Vector();
						v.type.id = LogicalTypeId::VARCHAR;
						v.type.physical_bytes = 16;
						v.data = lstate.filename_rows.data();
					} else { // union_by_name column this file does not have
						v = Vector();
						const idx_t g = column_ids[i];
						v.type = g < columns.size() ? columns[g].type : LogicalType();
						v.all_null = true;
					}
				}
				chunk.sel = file_chunk.sel;
				chunk.SetCardinality(file_chunk.count);
				return true;
			}
		}
		if (!TryInitializeNextBatch(gstate, lstate)) {
			chunk.SetCardinality(0);
			return false;
		}
	}
}

} // namespace flsreader
