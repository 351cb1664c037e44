// fls_writer.cpp — flsgpu_writer_*: row-group buffering + container assembly (host C++).
// Mirrors FastLanesFileWriter (reference src/writer/fls_writer.cpp): rows are buffered until a row group is
// full (:46-102), each column of a full row group goes through WriteColumn (:14-39), Finalize flushes the tail
// (:332-347).
#include "flsgpu.h"
#include "fls_encoder.hpp"
#include "fls_error.hpp"

#include <fcntl.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

using namespace flshost;

struct flsgpu_writer {
	std::vector<fls_column_desc> cols;
	std::vector<ColumnType> types;
	std::vector<int> force;
	uint32_t vec_per_rg = FLS_DEFAULT_VEC_PER_RG;
	uint64_t rg_tuples = 0;
	std::vector<std::vector<uint8_t>> pending; // per column, < rg_tuples rows (string columns: the bytes)
	std::vector<std::vector<uint64_t>> pending_off; // string columns: pending_rows + 1 offsets into pending[c]
	uint64_t pending_rows = 0;
	std::vector<fls_rowgroup_desc> rgs;
	std::vector<std::unique_ptr<EncodedChunk>> chunks; // row-group major
	uint64_t n_tuples = 0;
	bool finished = false;
	std::vector<uint8_t> image;
	bool detect_equal = true; // FLSGPU_WRITER_EQUAL=0: never write EXP_EQUAL chunks
	bool inline_footer = true; // WithInlinedFooter (fls_writer.cpp:278): false -> flsgpu_writer_save writes <path>.footer beside the data
	std::mutex flush_mu;       // flsgpu_writer_flush_rowgroup may be called from any thread
	// flsgpu_writer_stream_to: chunks go to the file as their row group is flushed / encoded and only their descriptors stay
	int fd = -1;
	std::string stream_path;
	uint64_t stream_off = FLS_ALIGN; // where the next chunk goes (the header takes the first 128 bytes)
	size_t streamed_chunks = 0;      // chunks[0 .. streamed_chunks) are in the file

	~flsgpu_writer() {
		if (fd >= 0) {
			::close(fd);
		}
	}
};

static int write_file(const char* path, const uint8_t* p, size_t n) {
	FILE* fp = std::fopen(path, "wb");
	if (!fp) {
		return fls_set_error(FLSGPU_ERR_IO, (std::string("cannot open for writing: ") + path).c_str());
	}
	const size_t k = std::fwrite(p, 1, n, fp);
	if (std::fclose(fp) != 0 || k != n) {
		return fls_set_error(FLSGPU_ERR_IO, (std::string("short write: ") + path).c_str());
	}
	return FLSGPU_OK;
}

static int pwrite_all(int fd, const uint8_t* p, size_t n, uint64_t off, const std::string& path) {
	while (n) {
		const ssize_t k = ::pwrite(fd, p, n, (off_t)off);
		if (k < 0) {
			if (errno == EINTR) {
				continue;
			}
			return fls_set_error(FLSGPU_ERR_IO, ("write failed: " + path + ": " + std::strerror(errno)).c_str());
		}
		p += k;
		n -= (size_t)k;
		off += (uint64_t)k;
	}
	return FLSGPU_OK;
}

// streaming writers: put the chunks that arrived since the last call into the file (same offsets as flsgpu_writer_finish lays
// an in-memory image out with) and drop their bytes
static int stream_new_chunks(flsgpu_writer* w) {
	if (w->fd < 0) {
		return FLSGPU_OK;
	}
	for (; w->streamed_chunks < w->chunks.size(); ++w->streamed_chunks) {
		EncodedChunk& ch = *w->chunks[w->streamed_chunks];
		if (ch.equal_to >= 0) {
			continue; // EXP_EQUAL: a descriptor only
		}
		ch.desc.offset = w->stream_off;
		const int rc = pwrite_all(w->fd, ch.bytes.data(), ch.bytes.size(), w->stream_off, w->stream_path);
		if (rc) {
			return rc;
		}
		w->stream_off += ch.desc.size;
		ch.bytes.clear();
		ch.bytes.shrink_to_fit();
	}
	return FLSGPU_OK;
}

// RowGroupWriter of the reference (fls_writer.cpp:306-330): one row group encoded off to the side by PrepareBatch, appended
// to the file by FlushBatch.
struct flsgpu_encoded_rowgroup {
	uint32_t n_tuples = 0;
	std::vector<std::unique_ptr<EncodedChunk>> chunks; // one per column
};

// Where the rows of one column start: fixed-width values, or (offsets, bytes) of a string column.
struct ColSrc {
	const uint8_t* data = nullptr;
	const uint64_t* offsets = nullptr; // string columns: offsets[i], offsets[i+1] delimit row i inside data
};

static bool is_string(const ColumnType& t) {
	return t.phys_bits == 128;
}

// One column chunk: rows [first_row, first_row + nt) of column c.  EXP_EQUAL (FastLanes multi-column compression; the reference
// detects its use in HasMccEncoding, fls_reader.cpp:43-82): a numeric column whose row group repeats an earlier column of the
// same physical type is stored as a reference.
static std::unique_ptr<EncodedChunk> encode_one(const flsgpu_writer* w, const std::vector<ColSrc>& col_ptrs, uint64_t first_row, uint32_t nt,
                                                size_t c, std::string* err_out) {
	const size_t esz = w->types[c].phys_bits / 8;
	auto ch = std::make_unique<EncodedChunk>();
	const char* err = "";
	if (w->detect_equal && !is_string(w->types[c]) && w->force[c] < 0) {
		const uint8_t* mine = col_ptrs[c].data + first_row * esz;
		for (uint64_t a = 0; a < c; ++a) {
			const ColumnType &ta = w->types[a], &tc = w->types[c];
			if (ta.phys_bits == tc.phys_bits && ta.is_signed == tc.is_signed && ta.is_float == tc.is_float &&
			    std::memcmp(col_ptrs[a].data + first_row * esz, mine, (size_t)nt * esz) == 0) {
				ch->equal_to = (int)a;
				return ch;
			}
		}
	}
	const bool ok = is_string(w->types[c]) ? encode_string_chunk(col_ptrs[c].offsets + first_row, col_ptrs[c].data, nt, w->force[c], *ch, &err)
	                                       : encode_chunk(w->types[c], col_ptrs[c].data + first_row * esz, nt, w->force[c], *ch, &err);
	if (!ok) {
		*err_out = std::string("column '") + w->cols[c].name + "': " + err;
		return nullptr;
	}
	return ch;
}

static int encode_rowgroups(flsgpu_writer* w, const std::vector<ColSrc>& col_ptrs, uint64_t n_rows, int n_threads) {
	// n_rows covers k full row groups, or one partial tail row group
	const uint64_t n_rg = (n_rows + w->rg_tuples - 1) / w->rg_tuples;
	const size_t n_cols = w->cols.size();
	const size_t first_chunk = w->chunks.size();
	w->chunks.resize(first_chunk + n_rg * n_cols);
	for (uint64_t r = 0; r < n_rg; ++r) {
		fls_rowgroup_desc rd {};
		rd.first_tuple = w->n_tuples + r * w->rg_tuples;
		rd.n_tuples = (uint32_t)std::min<uint64_t>(w->rg_tuples, n_rows - r * w->rg_tuples);
		rd.n_vec = (rd.n_tuples + FLS_VEC_SZ - 1) / FLS_VEC_SZ;
		w->rgs.push_back(rd);
	}
	std::atomic<uint64_t> next {0};
	std::atomic<bool> failed {false};
	std::string fail_msg;
	std::mutex* mu = new std::mutex();
	auto work = [&]() {
		for (;;) {
			const uint64_t task = next.fetch_add(1);
			if (task >= n_rg * n_cols || failed.load()) {
				return;
			}
			const uint64_t r = task / n_cols, c = task % n_cols;
			const uint32_t nt = (uint32_t)std::min<uint64_t>(w->rg_tuples, n_rows - r * w->rg_tuples);
			std::string err;
			auto ch = encode_one(w, col_ptrs, r * w->rg_tuples, nt, (size_t)c, &err);
			if (!ch) {
				std::lock_guard<std::mutex> g(*mu);
				failed = true;
				fail_msg = err;
				return;
			}
			w->chunks[first_chunk + task] = std::move(ch);
		}
	};
	const int nt = std::max(1, std::min<int>(n_threads, (int)(n_rg * n_cols)));
	if (nt == 1) {
		work();
	} else {
		std::vector<std::thread> th;
		for (int i = 0; i < nt; ++i) {
			th.emplace_back(work);
		}
		for (auto& t : th) {
			t.join();
		}
	}
	delete mu;
	if (failed) {
		w->chunks.resize(first_chunk);
		w->rgs.resize(w->rgs.size() - n_rg);
		return fls_set_error(FLSGPU_ERR_INVALID, fail_msg.c_str());
	}
	w->n_tuples += n_rows;
	return FLSGPU_OK;
}

extern "C" {

int flsgpu_writer_create(uint32_t n_columns, const flsgpu_column_spec* cols, uint32_t vec_per_rowgroup, flsgpu_writer** out) {
	if (!out || !cols || n_columns == 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_create: null argument or zero columns");
	}
	auto w = std::make_unique<flsgpu_writer>();
	w->vec_per_rg = vec_per_rowgroup ? vec_per_rowgroup : FLS_DEFAULT_VEC_PER_RG;
	if (const char* eq = std::getenv("FLSGPU_WRITER_EQUAL")) {
		w->detect_equal = std::atoi(eq) != 0;
	}
	w->rg_tuples = (uint64_t)w->vec_per_rg * FLS_VEC_SZ;
	for (uint32_t i = 0; i < n_columns; ++i) {
		const ColumnType t = column_type_of(cols[i].data_type, cols[i].decimal_precision);
		if (!t.phys_bits) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_writer_create: unsupported column data type");
		}
		fls_column_desc d {};
		std::snprintf(d.name, sizeof(d.name), "%s", cols[i].name ? cols[i].name : "");
		d.data_type = t.data_type;
		d.phys_bits = t.phys_bits;
		d.is_signed = t.is_signed;
		d.is_float = t.is_float;
		d.decimal_precision = (uint8_t)cols[i].decimal_precision;
		d.decimal_scale = (uint8_t)cols[i].decimal_scale;
		w->cols.push_back(d);
		w->types.push_back(t);
		w->force.push_back(cols[i].force_encoding);
	}
	w->pending.resize(n_columns);
	w->pending_off.resize(n_columns);
	for (uint32_t i = 0; i < n_columns; ++i) {
		if (is_string(w->types[i])) {
			w->pending_off[i].assign(1, 0);
		}
	}
	*out = w.release();
	return FLSGPU_OK;
}

int flsgpu_writer_append(flsgpu_writer* w, const void* const* columns, uint64_t n_tuples, int n_threads) {
	if (!w || !columns) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_append: null argument");
	}
	if (w->finished) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_append: writer already finished");
	}
	const size_t n_cols = w->cols.size();
	uint64_t consumed = 0;
	// rows [first, first + n) of the caller's column c appended to the pending buffer
	auto keep = [&](size_t c, uint64_t first, uint64_t n) {
		if (is_string(w->types[c])) {
			const flsgpu_string_column* sc = static_cast<const flsgpu_string_column*>(columns[c]);
			const uint64_t b0 = sc->offsets[first], b1 = sc->offsets[first + n];
			const uint64_t shift = w->pending[c].size() - b0;
			w->pending[c].insert(w->pending[c].end(), sc->bytes + b0, sc->bytes + b1);
			for (uint64_t i = 1; i <= n; ++i) {
				w->pending_off[c].push_back(sc->offsets[first + i] + shift);
			}
		} else {
			const size_t esz = w->types[c].phys_bits / 8;
			const uint8_t* src = static_cast<const uint8_t*>(columns[c]) + first * esz;
			w->pending[c].insert(w->pending[c].end(), src, src + n * esz);
		}
	};
	auto pending_src = [&]() {
		std::vector<ColSrc> ptrs(n_cols);
		for (size_t c = 0; c < n_cols; ++c) {
			ptrs[c].data = w->pending[c].data();
			ptrs[c].offsets = is_string(w->types[c]) ? w->pending_off[c].data() : nullptr;
		}
		return ptrs;
	};
	auto clear_pending = [&]() {
		for (size_t c = 0; c < n_cols; ++c) {
			w->pending[c].clear();
			if (is_string(w->types[c])) {
				w->pending_off[c].assign(1, 0);
			}
		}
		w->pending_rows = 0;
	};
	for (size_t c = 0; c < n_cols; ++c) {
		if (!columns[c]) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_append: null column pointer");
		}
	}
	// 1. top up the pending buffer to a full row group
	if (w->pending_rows) {
		const uint64_t take = std::min<uint64_t>(n_tuples, w->rg_tuples - w->pending_rows);
		for (size_t c = 0; c < n_cols; ++c) {
			keep(c, 0, take);
		}
		w->pending_rows += take;
		consumed = take;
		if (w->pending_rows == w->rg_tuples) {
			int rc = encode_rowgroups(w, pending_src(), w->rg_tuples, n_threads);
			if (!rc) {
				rc = stream_new_chunks(w);
			}
			if (rc) {
				return rc;
			}
			clear_pending();
		}
	}
	// 2. encode full row groups straight from the caller's arrays
	const uint64_t full = (n_tuples - consumed) / w->rg_tuples * w->rg_tuples;
	if (full) {
		std::vector<ColSrc> ptrs(n_cols);
		for (size_t c = 0; c < n_cols; ++c) {
			if (is_string(w->types[c])) {
				const flsgpu_string_column* sc = static_cast<const flsgpu_string_column*>(columns[c]);
				ptrs[c].data = sc->bytes;
				ptrs[c].offsets = sc->offsets + consumed;
			} else {
				ptrs[c].data = static_cast<const uint8_t*>(columns[c]) + consumed * (w->types[c].phys_bits / 8);
			}
		}
		int rc = encode_rowgroups(w, ptrs, full, n_threads);
		if (!rc) {
			rc = stream_new_chunks(w);
		}
		if (rc) {
			return rc;
		}
		consumed += full;
	}
	// 3. keep the remainder
	if (consumed < n_tuples) {
		const uint64_t rest = n_tuples - consumed;
		for (size_t c = 0; c < n_cols; ++c) {
			keep(c, consumed, rest);
		}
		w->pending_rows += rest;
	}
	return FLSGPU_OK;
}

// ---- row groups prepared out of order, flushed in order (PrepareBatch / FlushBatch, reference fls_writer.cpp:306-330) ----
int flsgpu_writer_encode_rowgroup(const flsgpu_writer* w, const void* const* columns, uint64_t n_tuples, flsgpu_encoded_rowgroup** out) {
	if (!w || !columns || !out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_encode_rowgroup: null argument");
	}
	if (n_tuples == 0 || n_tuples > w->rg_tuples) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_encode_rowgroup: a row group holds 1..row_group_size tuples");
	}
	const size_t n_cols = w->cols.size();
	std::vector<ColSrc> ptrs(n_cols);
	for (size_t c = 0; c < n_cols; ++c) {
		if (!columns[c]) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_encode_rowgroup: null column pointer");
		}
		if (is_string(w->types[c])) {
			const flsgpu_string_column* sc = static_cast<const flsgpu_string_column*>(columns[c]);
			ptrs[c].data = sc->bytes;
			ptrs[c].offsets = sc->offsets;
		} else {
			ptrs[c].data = static_cast<const uint8_t*>(columns[c]);
		}
	}
	auto rg = std::make_unique<flsgpu_encoded_rowgroup>();
	rg->n_tuples = (uint32_t)n_tuples;
	for (size_t c = 0; c < n_cols; ++c) {
		std::string err;
		auto ch = encode_one(w, ptrs, 0, (uint32_t)n_tuples, c, &err);
		if (!ch) {
			return fls_set_error(FLSGPU_ERR_INVALID, err.c_str());
		}
		rg->chunks.push_back(std::move(ch));
	}
	*out = rg.release();
	return FLSGPU_OK;
}

int flsgpu_writer_flush_rowgroup(flsgpu_writer* w, flsgpu_encoded_rowgroup* rg) {
	if (!w || !rg) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_flush_rowgroup: null argument");
	}
	std::unique_ptr<flsgpu_encoded_rowgroup> own(rg);
	std::lock_guard<std::mutex> g(w->flush_mu);
	if (w->finished || w->pending_rows || rg->chunks.size() != w->cols.size()) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_flush_rowgroup: writer finished, rows pending from flsgpu_writer_append, or foreign row group");
	}
	fls_rowgroup_desc rd {};
	rd.first_tuple = w->n_tuples;
	rd.n_tuples = rg->n_tuples;
	rd.n_vec = (rg->n_tuples + FLS_VEC_SZ - 1) / FLS_VEC_SZ;
	w->rgs.push_back(rd);
	for (auto& ch : rg->chunks) {
		w->chunks.push_back(std::move(ch));
	}
	w->n_tuples += rg->n_tuples;
	return stream_new_chunks(w);
}

int flsgpu_writer_stream_to(flsgpu_writer* w, const char* path) {
	if (!w || !path) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_stream_to: null argument");
	}
	std::lock_guard<std::mutex> g(w->flush_mu);
	if (w->finished || w->fd >= 0 || !w->chunks.empty()) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_stream_to: call once, before the first row group");
	}
	const int fd = ::open(path, O_CREAT | O_TRUNC | O_WRONLY | O_CLOEXEC, 0644);
	if (fd < 0) {
		return fls_set_error(FLSGPU_ERR_IO, (std::string("cannot open for writing: ") + path + ": " + std::strerror(errno)).c_str());
	}
	w->fd = fd;
	w->stream_path = path;
	return FLSGPU_OK;
}

void flsgpu_encoded_rowgroup_free(flsgpu_encoded_rowgroup* rg) {
	delete rg;
}

int flsgpu_writer_set_inline_footer(flsgpu_writer* w, int inline_footer) {
	if (!w || w->finished) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_set_inline_footer: null or finished writer");
	}
	w->inline_footer = inline_footer != 0;
	return FLSGPU_OK;
}

int flsgpu_writer_rowgroups(const flsgpu_writer* w, uint64_t* n_rowgroups, uint64_t* n_tuples) {
	if (!w) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_rowgroups: null writer");
	}
	if (n_rowgroups) {
		*n_rowgroups = w->rgs.size();
	}
	if (n_tuples) {
		*n_tuples = w->n_tuples + w->pending_rows;
	}
	return FLSGPU_OK;
}

int flsgpu_writer_finish(flsgpu_writer* w) {
	if (!w) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_finish: null writer");
	}
	if (w->finished) {
		return FLSGPU_OK;
	}
	if (w->pending_rows) {
		std::vector<ColSrc> ptrs(w->cols.size());
		for (size_t c = 0; c < w->cols.size(); ++c) {
			ptrs[c].data = w->pending[c].data();
			ptrs[c].offsets = is_string(w->types[c]) ? w->pending_off[c].data() : nullptr;
		}
		int rc = encode_rowgroups(w, ptrs, w->pending_rows, 1);
		if (!rc) {
			rc = stream_new_chunks(w);
		}
		if (rc) {
			return rc;
		}
		w->pending_rows = 0;
		for (auto& p : w->pending) {
			p.clear();
			p.shrink_to_fit();
		}
	}
	// lay out: header | chunks (128-aligned) | footer
	uint64_t off = FLS_ALIGN; // header occupies the first 128 bytes
	const size_t n_cols_all = w->cols.size();
	for (auto& ch : w->chunks) {
		if (ch->equal_to < 0) {
			ch->desc.offset = off;
			off += ch->desc.size;
		}
	}
	for (size_t i = 0; i < w->chunks.size(); ++i) { // EXP_EQUAL chunks: the descriptor of the column they repeat (never itself one)
		EncodedChunk& ch = *w->chunks[i];
		if (ch.equal_to >= 0) {
			ch.desc = w->chunks[i - i % n_cols_all + (size_t)ch.equal_to]->desc;
			ch.desc.is_equal = 1;
			ch.desc.equal_col[0] = (uint8_t)(ch.equal_to & 0xff);
			ch.desc.equal_col[1] = (uint8_t)((ch.equal_to >> 8) & 0xff);
		}
	}
	const uint64_t footer_off = off;
	const uint64_t footer_size = w->cols.size() * sizeof(fls_column_desc) + w->rgs.size() * sizeof(fls_rowgroup_desc) +
	                             w->chunks.size() * sizeof(fls_chunk_desc);
	fls_file_header h {};
	std::memcpy(h.magic, FLS_MAGIC, 8);
	h.version = FLS_VERSION;
	h.n_columns = (uint32_t)w->cols.size();
	h.n_rowgroups = (uint32_t)w->rgs.size();
	h.vec_per_rowgroup = w->vec_per_rg;
	h.n_tuples = w->n_tuples;
	h.footer_offset = footer_off;
	h.footer_size = footer_size;
	h.reserved[0] = w->inline_footer ? 0 : 1; // 1: flsgpu_writer_save puts the footer into <path>.footer
	const bool streaming = w->fd >= 0;
	// streaming writers hold the header + footer only: the chunks are in the file already, at these very offsets
	std::vector<uint8_t> tail;
	uint8_t* f;
	if (streaming) {
		if (w->streamed_chunks != w->chunks.size() || w->stream_off != footer_off) {
			return fls_set_error(FLSGPU_ERR_IO, "flsgpu_writer_finish: streamed chunks disagree with the final layout");
		}
		tail.assign(footer_size, 0);
		f = tail.data();
	} else {
		w->image.assign(footer_off + footer_size, 0);
		std::memcpy(w->image.data(), &h, sizeof(h));
		f = w->image.data() + footer_off;
	}
	std::memcpy(f, w->cols.data(), w->cols.size() * sizeof(fls_column_desc));
	f += w->cols.size() * sizeof(fls_column_desc);
	if (!w->rgs.empty()) { // an empty table has columns only
		std::memcpy(f, w->rgs.data(), w->rgs.size() * sizeof(fls_rowgroup_desc));
	}
	f += w->rgs.size() * sizeof(fls_rowgroup_desc);
	for (auto& ch : w->chunks) {
		if (!streaming && !ch->bytes.empty()) { // EXP_EQUAL chunks have a descriptor only
			std::memcpy(w->image.data() + ch->desc.offset, ch->bytes.data(), ch->bytes.size());
		}
		std::memcpy(f, &ch->desc, sizeof(fls_chunk_desc));
		f += sizeof(fls_chunk_desc);
		ch->bytes.clear();
		ch->bytes.shrink_to_fit();
	}
	if (streaming) {
		uint8_t head[FLS_ALIGN] = {};
		std::memcpy(head, &h, sizeof(h));
		int rc = pwrite_all(w->fd, head, sizeof(head), 0, w->stream_path);
		if (!rc && ::ftruncate(w->fd, (off_t)(w->inline_footer ? footer_off + footer_size : footer_off)) != 0) {
			rc = fls_set_error(FLSGPU_ERR_IO, ("cannot size " + w->stream_path + ": " + std::strerror(errno)).c_str());
		}
		if (!rc && w->inline_footer) {
			rc = pwrite_all(w->fd, tail.data(), tail.size(), footer_off, w->stream_path);
		}
		if (::close(w->fd) != 0 && !rc) {
			rc = fls_set_error(FLSGPU_ERR_IO, ("close failed: " + w->stream_path).c_str());
		}
		w->fd = -1;
		if (!rc && !w->inline_footer) {
			rc = write_file((w->stream_path + ".footer").c_str(), tail.data(), tail.size());
		}
		if (rc) {
			return rc;
		}
	}
	w->finished = true;
	return FLSGPU_OK;
}

int flsgpu_writer_image(const flsgpu_writer* w, const void** bytes, uint64_t* size) {
	if (!w || !w->finished || !bytes || !size) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_image: writer not finished");
	}
	if (!w->stream_path.empty()) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_image: a streaming writer keeps no image (flsgpu_writer_stream_to)");
	}
	*bytes = w->image.data();
	*size = w->image.size();
	return FLSGPU_OK;
}

int flsgpu_writer_save(const flsgpu_writer* w, const char* path) {
	if (!w || !w->finished || !path) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_save: writer not finished");
	}
	if (!w->stream_path.empty()) { // the file was complete when flsgpu_writer_finish returned
		return w->stream_path == path ? FLSGPU_OK
		                              : fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_writer_save: a streaming writer saves to the path it streams to");
	}
	if (w->inline_footer) {
		return write_file(path, w->image.data(), w->image.size());
	}
	// WithInlinedFooter(false) (reference fls_writer.cpp:229-230,278): data file = header + chunks, footer beside it
	fls_file_header h;
	std::memcpy(&h, w->image.data(), sizeof(h));
	const int rc = write_file(path, w->image.data(), (size_t)h.footer_offset);
	if (rc) {
		return rc;
	}
	return write_file((std::string(path) + ".footer").c_str(), w->image.data() + h.footer_offset, (size_t)h.footer_size);
}

void flsgpu_writer_destroy(flsgpu_writer* w) {
	delete w;
}

} // extern "C"
