// fls_copy_function.cpp — COPY ... TO 'x.fls' (FORMAT fls) over the engine's host-side writer (flsgpu_writer_*).
//
// Same callbacks, options and row-group discipline as the reference writer (src/write_fls.cpp:8-29,
// src/writer/fls_writer.cpp): rows are buffered per thread until a row group (row_group_size tuples, a multiple of 1024) is
// full, the row group is encoded off to the side and appended to the file; thread remainders meet in a combine buffer;
// Finalize flushes the tail row group.  Batch mode prepares row groups out of order and flushes them in order.
// The encodings themselves are chosen by the engine's encoder (in the reference: by the absent FastLanes library).
#include "fastlanes_extension.hpp"
#include "fls_file_reader.hpp"

#include "duckdb/common/string_util.hpp"
#include "duckdb/common/types/column/column_data_collection.hpp"
#include "duckdb/common/types/decimal.hpp"
#include "duckdb/common/types/hugeint.hpp"
#include "duckdb/common/types/uhugeint.hpp"
#include "duckdb/function/copy_function.hpp"
#include "duckdb/main/extension/extension_loader.hpp"

#include <cstring>
#include <mutex>

namespace duckdb {

namespace {

// ---- type translation (reference src/writer/translation_utils.cpp:6-70, fls_view_writer.cpp:61-94) -----------------
enum class RowFormat : uint8_t { FIXED, STRING, HUGEINT_TEXT, UHUGEINT_TEXT, DECIMAL128_TEXT };

struct ColumnPlan {
	flsgpu_column_spec spec {};
	RowFormat format = RowFormat::FIXED;
	idx_t width = 0; // FIXED: bytes per value
};

ColumnPlan PlanColumn(const string &name, const LogicalType &type) {
	ColumnPlan plan;
	plan.spec.name = name.c_str();
	plan.spec.force_encoding = -1;
	auto fixed = [&](int32_t data_type, idx_t width) {
		plan.spec.data_type = data_type;
		plan.width = width;
	};
	auto text = [&](int32_t data_type, RowFormat format) {
		plan.spec.data_type = data_type;
		plan.format = format;
	};
	switch (type.id()) {
	case LogicalTypeId::BOOLEAN:
		fixed(FLS_TYPE_BOOLEAN, 1);
		break;
	case LogicalTypeId::TINYINT:
		fixed(FLS_TYPE_INT8, 1);
		break;
	case LogicalTypeId::SMALLINT:
		fixed(FLS_TYPE_INT16, 2);
		break;
	case LogicalTypeId::INTEGER:
		fixed(FLS_TYPE_INT32, 4);
		break;
	case LogicalTypeId::BIGINT:
		fixed(FLS_TYPE_INT64, 8);
		break;
	case LogicalTypeId::UTINYINT:
		fixed(FLS_TYPE_UINT8, 1);
		break;
	case LogicalTypeId::USMALLINT:
		fixed(FLS_TYPE_UINT16, 2);
		break;
	case LogicalTypeId::UINTEGER:
		fixed(FLS_TYPE_UINT32, 4);
		break;
	case LogicalTypeId::UBIGINT:
		fixed(FLS_TYPE_UINT64, 8);
		break;
	case LogicalTypeId::FLOAT:
		fixed(FLS_TYPE_FLOAT, 4);
		break;
	case LogicalTypeId::DOUBLE:
		fixed(FLS_TYPE_DOUBLE, 8);
		break;
	case LogicalTypeId::DATE:
		fixed(FLS_TYPE_DATE, 4);
		break;
	case LogicalTypeId::TIMESTAMP:
		fixed(FLS_TYPE_TIMESTAMP, 8);
		break;
	case LogicalTypeId::DECIMAL:
		// physical INT16/32/64 by width, annotated with (precision, scale) so that the reader restores the DECIMAL type
		// (fls_writer.cpp:258-266); 128-bit decimals travel as decimal strings (translation_utils.cpp:32-33)
		if (type.InternalType() == PhysicalType::INT128) {
			text(FLS_TYPE_FLS_STR, RowFormat::DECIMAL128_TEXT);
		} else {
			fixed(FLS_TYPE_DECIMAL, GetTypeIdSize(type.InternalType()));
			plan.spec.decimal_precision = DecimalType::GetWidth(type);
			plan.spec.decimal_scale = DecimalType::GetScale(type);
		}
		break;
	case LogicalTypeId::HUGEINT:
		text(FLS_TYPE_FLS_STR, RowFormat::HUGEINT_TEXT);
		break;
	case LogicalTypeId::UHUGEINT:
		text(FLS_TYPE_FLS_STR, RowFormat::UHUGEINT_TEXT);
		break;
	case LogicalTypeId::VARCHAR:
	case LogicalTypeId::CHAR:
		text(FLS_TYPE_FLS_STR, RowFormat::STRING);
		break;
	case LogicalTypeId::BLOB:
		text(FLS_TYPE_BYTE_ARRAY, RowFormat::STRING);
		break;
	case LogicalTypeId::LIST:
	case LogicalTypeId::STRUCT:
	case LogicalTypeId::MAP:
		throw NotImplementedException("PhysicalType %s is not supported by FastLanes writer", TypeIdToString(type.InternalType()));
	default:
		throw InvalidInputException("FastLanesWriter: column has unsupported type \"%s\"", type.ToString());
	}
	return plan;
}

// ---- row-group buffer: the rows of one row group in the layout flsgpu_writer_encode_rowgroup takes -------------------
struct RowGroupBuffer {
	explicit RowGroupBuffer(const vector<ColumnPlan> &plans_p) : plans(plans_p), fixed(plans_p.size()), offsets(plans_p.size()), bytes(plans_p.size()) {
		Reset();
	}

	const vector<ColumnPlan> &plans;
	vector<vector<data_t>> fixed;
	vector<vector<uint64_t>> offsets;
	vector<vector<data_t>> bytes;
	idx_t count = 0;

	void Reset() {
		count = 0;
		for (idx_t c = 0; c < plans.size(); c++) {
			fixed[c].clear();
			bytes[c].clear();
			offsets[c].assign(plans[c].format == RowFormat::FIXED ? 0 : 1, 0);
		}
	}

	void AppendText(idx_t c, const char *data, idx_t size) {
		bytes[c].insert(bytes[c].end(), const_data_ptr_cast(data), const_data_ptr_cast(data) + size);
		offsets[c].push_back(bytes[c].size());
	}

	//! Rows [first, first + n) of the chunk.  FastLanes has no NULLs (the reference writes whatever the flat vector holds,
	//! fls_view_writer.hpp:20-25); NULL rows are written as zero / the empty string here.
	void Append(DataChunk &chunk, idx_t first, idx_t n) {
		for (idx_t c = 0; c < plans.size(); c++) {
			auto &plan = plans[c];
			UnifiedVectorFormat format;
			chunk.data[c].ToUnifiedFormat(chunk.size(), format);
			const bool dense = !format.sel->IsSet() && format.validity.AllValid();
			if (plan.format == RowFormat::FIXED) {
				const idx_t at = fixed[c].size();
				fixed[c].resize(at + n * plan.width);
				if (dense) {
					std::memcpy(fixed[c].data() + at, format.data + first * plan.width, n * plan.width);
					continue;
				}
				for (idx_t i = 0; i < n; i++) {
					const auto idx = format.sel->get_index(first + i);
					if (format.validity.RowIsValid(idx)) {
						std::memcpy(fixed[c].data() + at + i * plan.width, format.data + idx * plan.width, plan.width);
					} else {
						std::memset(fixed[c].data() + at + i * plan.width, 0, plan.width);
					}
				}
				continue;
			}
			for (idx_t i = 0; i < n; i++) {
				const auto idx = format.sel->get_index(first + i);
				if (!format.validity.RowIsValid(idx)) {
					AppendText(c, "", 0);
					continue;
				}
				switch (plan.format) {
				case RowFormat::STRING: {
					const auto &s = UnifiedVectorFormat::GetData<string_t>(format)[idx];
					AppendText(c, s.GetData(), s.GetSize());
					break;
				}
				case RowFormat::HUGEINT_TEXT: { // fls_view_writer.cpp:28-36
					const auto s = UnifiedVectorFormat::GetData<hugeint_t>(format)[idx].ToString();
					AppendText(c, s.data(), s.size());
					break;
				}
				case RowFormat::UHUGEINT_TEXT: { // fls_view_writer.cpp:46-54
					const auto s = UnifiedVectorFormat::GetData<uhugeint_t>(format)[idx].ToString();
					AppendText(c, s.data(), s.size());
					break;
				}
				default: {
					auto &type = chunk.data[c].GetType();
					const auto s = Decimal::ToString(UnifiedVectorFormat::GetData<hugeint_t>(format)[idx], DecimalType::GetWidth(type),
					                                 DecimalType::GetScale(type));
					AppendText(c, s.data(), s.size());
					break;
				}
				}
			}
		}
		count += n;
	}

	void Append(RowGroupBuffer &other, idx_t first, idx_t n) {
		for (idx_t c = 0; c < plans.size(); c++) {
			if (plans[c].format == RowFormat::FIXED) {
				const idx_t w = plans[c].width;
				fixed[c].insert(fixed[c].end(), other.fixed[c].begin() + first * w, other.fixed[c].begin() + (first + n) * w);
				continue;
			}
			const uint64_t b0 = other.offsets[c][first], b1 = other.offsets[c][first + n];
			const uint64_t shift = bytes[c].size() - b0;
			bytes[c].insert(bytes[c].end(), other.bytes[c].begin() + b0, other.bytes[c].begin() + b1);
			for (idx_t i = 1; i <= n; i++) {
				offsets[c].push_back(other.offsets[c][first + i] + shift);
			}
		}
		count += n;
	}

	//! RowGroupWriter: WriteColumn for every column + Finalize (fls_writer.cpp:14-39)
	flsgpu_encoded_rowgroup *Encode(flsgpu_writer *writer) const {
		vector<flsgpu_string_column> string_columns(plans.size());
		vector<const void *> column_ptrs(plans.size());
		for (idx_t c = 0; c < plans.size(); c++) {
			if (plans[c].format == RowFormat::FIXED) {
				column_ptrs[c] = fixed[c].data();
			} else {
				string_columns[c].offsets = offsets[c].data();
				string_columns[c].bytes = bytes[c].empty() ? const_data_ptr_cast("") : bytes[c].data();
				column_ptrs[c] = &string_columns[c];
			}
		}
		flsgpu_encoded_rowgroup *encoded = nullptr;
		FlsCheck(flsgpu_writer_encode_rowgroup(writer, column_ptrs.data(), count, &encoded), "COPY (FORMAT fls)");
		return encoded;
	}
};

struct FastLanesWriteBindData : public TableFunctionData {
	//! Tuples per row group; a multiple of 1024 (reference fls_writer.hpp:13-14: 64 vectors x 1024)
	uint64_t row_group_size = uint64_t(FLS_DEFAULT_VEC_PER_RG) * FLS_VEC_SZ;
	//! Rotate to a new file after this many row groups (0: never)
	idx_t row_groups_per_file = 0;
	//! false: the footer goes to "<file>.footer"
	bool inline_footer = true;
	vector<LogicalType> types;
	vector<string> names;
	vector<ColumnPlan> plans;
};

struct FastLanesWriteGlobalState : public GlobalFunctionData {
	~FastLanesWriteGlobalState() override {
		flsgpu_writer_destroy(writer);
	}
	string file_path;
	flsgpu_writer *writer = nullptr;
	std::mutex combine_lock;
	unique_ptr<RowGroupBuffer> combine_buffer;
	atomic<idx_t> num_row_groups {0};

	//! batch mode: PrepareBatch runs on any thread and has no local state; its row-group buffers are recycled here so that
	//! every batch does not fault a few megabytes of fresh pages in
	std::mutex pool_lock;
	vector<unique_ptr<RowGroupBuffer>> pool;

	unique_ptr<RowGroupBuffer> AcquireBuffer(const vector<ColumnPlan> &plans) {
		{
			std::lock_guard<std::mutex> guard(pool_lock);
			if (!pool.empty()) {
				auto buffer = std::move(pool.back());
				pool.pop_back();
				return buffer;
			}
		}
		return make_uniq<RowGroupBuffer>(plans);
	}
	void ReleaseBuffer(unique_ptr<RowGroupBuffer> buffer) {
		buffer->Reset();
		std::lock_guard<std::mutex> guard(pool_lock);
		pool.push_back(std::move(buffer));
	}

	//! rg_writer->Flush() (fls_writer.cpp:70,94,130,161,325)
	void Flush(flsgpu_encoded_rowgroup *encoded) {
		FlsCheck(flsgpu_writer_flush_rowgroup(writer, encoded), "COPY (FORMAT fls)");
		num_row_groups++;
	}
};

struct FastLanesWriteLocalState : public LocalFunctionData {
	explicit FastLanesWriteLocalState(const vector<ColumnPlan> &plans) : buffer(plans) {
	}
	RowGroupBuffer buffer;
};

struct FastLanesWriteBatchData : public PreparedBatchData {
	~FastLanesWriteBatchData() override {
		for (auto rg : row_groups) {
			flsgpu_encoded_rowgroup_free(rg);
		}
	}
	vector<flsgpu_encoded_rowgroup *> row_groups;
};

// ---- callbacks -------------------------------------------------------------------------------------------------------

unique_ptr<FunctionData> FlsWriteBind(ClientContext &, CopyFunctionBindInput &input, const vector<string> &names,
                                      const vector<LogicalType> &sql_types) {
	auto data = make_uniq<FastLanesWriteBindData>();
	for (const auto &option : input.info.options) {
		const auto key = StringUtil::Lower(option.first);
		if (option.second.size() != 1) {
			throw BinderException("Only one value allowed per option.");
		}
		if (key == "row_group_size") {
			data->row_group_size = option.second[0].GetValue<uint64_t>();
		} else if (key == "row_groups_per_file") {
			data->row_groups_per_file = option.second[0].GetValue<uint64_t>();
		} else if (key == "inline_footer") {
			data->inline_footer = option.second[0].GetValue<bool>();
		} else {
			throw BinderException("Unrecognized option: " + option.first);
		}
	}
	if (data->row_group_size == 0 || data->row_group_size % FLS_VEC_SZ != 0) {
		throw BinderException("row_group_size must be a positive multiple of 1024 (the FastLanes vector size)");
	}
	data->types = sql_types;
	data->names = names;
	for (idx_t i = 0; i < names.size(); i++) {
		data->plans.push_back(PlanColumn(data->names[i], sql_types[i]));
	}
	return std::move(data);
}

unique_ptr<GlobalFunctionData> FlsWriteInitGlobal(ClientContext &, FunctionData &bind_data_p, const string &file_path) {
	auto &bind_data = bind_data_p.Cast<FastLanesWriteBindData>();
	auto global_state = make_uniq<FastLanesWriteGlobalState>();
	global_state->file_path = file_path;
	vector<flsgpu_column_spec> specs;
	for (idx_t i = 0; i < bind_data.plans.size(); i++) {
		specs.push_back(bind_data.plans[i].spec);
		specs.back().name = bind_data.names[i].c_str();
	}
	FlsCheck(flsgpu_writer_create(NumericCast<uint32_t>(specs.size()), specs.data(), NumericCast<uint32_t>(bind_data.row_group_size / FLS_VEC_SZ),
	                              &global_state->writer),
	         "COPY (FORMAT fls)");
	FlsCheck(flsgpu_writer_set_inline_footer(global_state->writer, bind_data.inline_footer ? 1 : 0), "COPY (FORMAT fls)");
	// row groups go to the file as they are flushed (rg_writer->Flush(), fls_writer.cpp:70,94): the writer keeps descriptors only
	FlsCheck(flsgpu_writer_stream_to(global_state->writer, file_path.c_str()), "COPY (FORMAT fls)");
	return std::move(global_state);
}

unique_ptr<LocalFunctionData> FlsWriteInitLocal(ExecutionContext &, FunctionData &bind_data_p) {
	return make_uniq<FastLanesWriteLocalState>(bind_data_p.Cast<FastLanesWriteBindData>().plans);
}

// Sink (fls_writer.cpp:46-102): fill the thread's buffer; every time it holds a full row group, encode and flush it
void FlsWriteSink(ExecutionContext &, FunctionData &bind_data_p, GlobalFunctionData &global_state_p, LocalFunctionData &local_state_p,
                  DataChunk &input) {
	auto &bind_data = bind_data_p.Cast<FastLanesWriteBindData>();
	auto &global_state = global_state_p.Cast<FastLanesWriteGlobalState>();
	auto &buffer = local_state_p.Cast<FastLanesWriteLocalState>().buffer;
	idx_t consumed = 0;
	while (consumed < input.size()) {
		const idx_t take = MinValue<idx_t>(input.size() - consumed, bind_data.row_group_size - buffer.count);
		buffer.Append(input, consumed, take);
		consumed += take;
		if (buffer.count == bind_data.row_group_size) {
			global_state.Flush(buffer.Encode(global_state.writer));
			buffer.Reset();
		}
	}
}

// Combine (fls_writer.cpp:104-187): thread remainders are merged into row groups of the full size under a lock
void FlsWriteCombine(ExecutionContext &, FunctionData &bind_data_p, GlobalFunctionData &global_state_p,
                     LocalFunctionData &local_state_p) {
	auto &bind_data = bind_data_p.Cast<FastLanesWriteBindData>();
	auto &global_state = global_state_p.Cast<FastLanesWriteGlobalState>();
	auto &buffer = local_state_p.Cast<FastLanesWriteLocalState>().buffer;
	if (buffer.count == 0) {
		return;
	}
	std::lock_guard<std::mutex> guard(global_state.combine_lock);
	if (!global_state.combine_buffer) {
		global_state.combine_buffer = make_uniq<RowGroupBuffer>(bind_data.plans);
	}
	auto &combine = *global_state.combine_buffer;
	idx_t consumed = 0;
	while (consumed < buffer.count) {
		const idx_t take = MinValue<idx_t>(buffer.count - consumed, bind_data.row_group_size - combine.count);
		combine.Append(buffer, consumed, take);
		consumed += take;
		if (combine.count == bind_data.row_group_size) {
			global_state.Flush(combine.Encode(global_state.writer));
			combine.Reset();
		}
	}
	buffer.Reset();
}

// Finalize (fls_writer.cpp:332-347): the tail row group, then the footer
void FlsWriteFinalize(ClientContext &, FunctionData &, GlobalFunctionData &global_state_p) {
	auto &global_state = global_state_p.Cast<FastLanesWriteGlobalState>();
	if (global_state.combine_buffer && global_state.combine_buffer->count) {
		global_state.Flush(global_state.combine_buffer->Encode(global_state.writer));
		global_state.combine_buffer->Reset();
	}
	FlsCheck(flsgpu_writer_finish(global_state.writer), "COPY (FORMAT fls)");
	FlsCheck(flsgpu_writer_save(global_state.writer, global_state.file_path.c_str()), "COPY (FORMAT fls)");
}

CopyFunctionExecutionMode FlsWriteExecutionMode(bool preserve_insertion_order, bool supports_batch_index) {
	if (!preserve_insertion_order) {
		return CopyFunctionExecutionMode::PARALLEL_COPY_TO_FILE;
	}
	if (supports_batch_index) {
		return CopyFunctionExecutionMode::BATCH_COPY_TO_FILE;
	}
	return CopyFunctionExecutionMode::REGULAR_COPY_TO_FILE;
}

idx_t FlsWriteDesiredBatchSize(ClientContext &, FunctionData &bind_data_p) {
	return bind_data_p.Cast<FastLanesWriteBindData>().row_group_size;
}

// PrepareBatch (fls_writer.cpp:306-320): may run out of order on any thread; encodes, does not touch the file
unique_ptr<PreparedBatchData> FlsWritePrepareBatch(ClientContext &, FunctionData &bind_data_p, GlobalFunctionData &global_state_p,
                                                   unique_ptr<ColumnDataCollection> collection) {
	auto &bind_data = bind_data_p.Cast<FastLanesWriteBindData>();
	auto &global_state = global_state_p.Cast<FastLanesWriteGlobalState>();
	auto batch = make_uniq<FastLanesWriteBatchData>();
	auto buffer_p = global_state.AcquireBuffer(bind_data.plans);
	auto &buffer = *buffer_p;
	for (auto &chunk : collection->Chunks()) {
		idx_t consumed = 0;
		while (consumed < chunk.size()) {
			const idx_t take = MinValue<idx_t>(chunk.size() - consumed, bind_data.row_group_size - buffer.count);
			buffer.Append(chunk, consumed, take);
			consumed += take;
			if (buffer.count == bind_data.row_group_size) {
				batch->row_groups.push_back(buffer.Encode(global_state.writer));
				buffer.Reset();
			}
		}
	}
	if (buffer.count) {
		batch->row_groups.push_back(buffer.Encode(global_state.writer));
	}
	global_state.ReleaseBuffer(std::move(buffer_p));
	return std::move(batch);
}

// FlushBatch (fls_writer.cpp:322-330): called in batch order
void FlsWriteFlushBatch(ClientContext &, FunctionData &, GlobalFunctionData &global_state_p, PreparedBatchData &batch_p) {
	auto &global_state = global_state_p.Cast<FastLanesWriteGlobalState>();
	auto &batch = batch_p.Cast<FastLanesWriteBatchData>();
	for (auto &rg : batch.row_groups) {
		auto encoded = rg;
		rg = nullptr; // flush takes ownership, whether it succeeds or not
		global_state.Flush(encoded);
	}
	batch.row_groups.clear();
}

bool FlsWriteRotateFiles(FunctionData &bind_data_p, const optional_idx &) {
	return bind_data_p.Cast<FastLanesWriteBindData>().row_groups_per_file != 0;
}

bool FlsWriteRotateNextFile(GlobalFunctionData &global_state_p, FunctionData &bind_data_p, const optional_idx &) {
	auto &bind_data = bind_data_p.Cast<FastLanesWriteBindData>();
	return bind_data.row_groups_per_file && global_state_p.Cast<FastLanesWriteGlobalState>().num_row_groups >= bind_data.row_groups_per_file;
}

} // namespace

void FastLanesCopyFunction::Register(ExtensionLoader &loader) {
	CopyFunction fn("fls");
	fn.copy_to_bind = FlsWriteBind;
	fn.copy_to_initialize_global = FlsWriteInitGlobal;
	fn.copy_to_initialize_local = FlsWriteInitLocal;
	fn.copy_to_sink = FlsWriteSink;
	fn.copy_to_combine = FlsWriteCombine;
	fn.copy_to_finalize = FlsWriteFinalize;
	fn.execution_mode = FlsWriteExecutionMode;
	fn.prepare_batch = FlsWritePrepareBatch;
	fn.flush_batch = FlsWriteFlushBatch;
	fn.desired_batch_size = FlsWriteDesiredBatchSize;
	fn.rotate_files = FlsWriteRotateFiles;
	fn.rotate_next_file = FlsWriteRotateNextFile;
	fn.extension = "fls";
	loader.RegisterFunction(fn);
}

} // namespace duckdb
