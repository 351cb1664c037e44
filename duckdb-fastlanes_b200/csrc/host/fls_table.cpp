// fls_table.cpp — flsgpu_table_open*/info/close: container parsing and validation (host C++).
#include "flsgpu.h"
#include "fls_error.hpp"
#include "fls_table.hpp"
#include "fls_fsst.hpp"

#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <algorithm>
#include <atomic>
#include <unistd.h>
#include <chrono>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

static thread_local std::string g_last_error;

int fls_set_error(int status, const char* msg) {
	g_last_error = msg ? msg : "";
	return status;
}

static bool is_patched(unsigned enc) {
	return enc == FLS_ENC_FFOR || enc == FLS_ENC_SLPATCH || enc == FLS_ENC_FREQUENCY || enc == FLS_ENC_ALP || enc == FLS_ENC_ALP_RD;
}

// packed part of a vector's blob (the exception tail follows it)
static unsigned packed_bytes(const fls_chunk_desc& c, const fls_vec_meta& m) {
	switch (c.encoding) {
	case FLS_ENC_UNCOMPRESSED: return 1024u * c.phys_bits / 8;
	case FLS_ENC_DELTA:
	case FLS_ENC_RLE: return 128u + 128u * m.width;
	case FLS_ENC_ALP_RD: return 128u * (m.width + m.width2);
	case FLS_ENC_CONSTANT:
	case FLS_ENC_CROSS_RLE: return 0;
	default: return 128u * m.width;
	}
}
static unsigned blob_bytes(const fls_chunk_desc& c, const fls_vec_meta& m) {
	return packed_bytes(c, m) + (is_patched(c.encoding) ? FLS_EXC_TAIL_BYTES(m.aux_cnt, c.phys_bits / 8u) : 0u);
}

static bool is_string_enc(unsigned enc) {
	return enc == FLS_ENC_DICT_STR || enc == FLS_ENC_STR_UNCOMPRESSED || enc == FLS_ENC_STR_CONSTANT || FLS_ENC_IS_FSST_ANY(enc) ||
	       FLS_ENC_IS_STR_VALUES(enc);
}

// String chunk (fls_format.h "String columns"): every offset array must be monotone and end at its byte segment's size,
// packed codes must stay inside the dictionary's lane width, and FSST code streams must decode to exactly the lengths the
// decoded end offsets promise (the GPU writes each string at its promised heap position without further checks).
static const char* validate_string_chunk(const flsgpu_table* t, const fls_chunk_desc& ch) {
	const uint8_t* base = t->bytes + ch.offset;
	const uint8_t* aux0 = base + ch.seg_off[FLS_SEG_AUX0];
	const uint32_t* aux1 = reinterpret_cast<const uint32_t*>(base + ch.seg_off[FLS_SEG_AUX1]);
	const fls_vec_meta* meta = reinterpret_cast<const fls_vec_meta*>(base + ch.seg_off[FLS_SEG_META]);
	const bool dict = FLS_ENC_IS_STR_DICT(ch.encoding);
	const bool fsst = FLS_ENC_IS_FSST_ANY(ch.encoding), fsst12 = FLS_ENC_IS_FSST12(ch.encoding);
	for (uint32_t v = 0; v < ch.n_vec; ++v) {
		if (meta[v].n_valid == 0 || meta[v].n_valid > FLS_VEC_SZ) {
			return "vector n_valid out of range";
		}
	}
	if (ch.encoding == FLS_ENC_STR_CONSTANT) {
		return (ch.seg_size[FLS_SEG_AUX0] == ch.aux_count && ch.const_val == ch.aux_count) ? nullptr : "constant string segment has the wrong size";
	}
	if (FLS_ENC_IS_STR_VALUES(ch.encoding)) {
		// value strings: bytes in AUX0, monotone end offsets covering them in AUX1 (CROSS_RLE_STR: followed by the run lengths)
		const uint64_t nv = ch.aux_count;
		const bool cross = ch.encoding == FLS_ENC_CROSS_RLE_STR;
		if (nv == 0 || ch.seg_size[FLS_SEG_AUX1] != nv * 4 * (cross ? 2 : 1)) {
			return "value string offsets have the wrong size";
		}
		uint32_t prev = 0;
		for (uint64_t i = 0; i < nv; ++i) {
			if (aux1[i] < prev) {
				return "string end offsets are not monotone";
			}
			prev = aux1[i];
		}
		if (prev != ch.seg_size[FLS_SEG_AUX0] || ch.const_val != prev) {
			return "string end offsets do not cover the byte segment";
		}
		if (cross) {
			uint64_t sum = 0;
			for (uint64_t i = 0; i < nv; ++i) {
				if (aux1[nv + i] == 0) {
					return "CROSS_RLE_STR run of length zero";
				}
				sum += aux1[nv + i];
			}
			return sum == (uint64_t)ch.n_vec * FLS_VEC_SZ ? nullptr : "CROSS_RLE_STR run lengths do not cover the chunk";
		}
		const uint8_t* data_seg = base + ch.seg_off[FLS_SEG_DATA];
		uint64_t next = 0; // the vectors' value ranges follow each other and cover the chunk's values
		for (uint32_t v = 0; v < ch.n_vec; ++v) {
			const fls_vec_meta& m = meta[v];
			const bool rle = ch.encoding == FLS_ENC_RLE_STR;
			const uint64_t n_vals = rle ? m.aux_cnt : (uint64_t)m.aux_cnt + 1u;
			const uint64_t blob = rle ? 128u + 128u * m.width : FLS_PAD16((uint64_t)m.aux_cnt * 2u);
			if (m.aux_off != next || n_vals == 0 || m.aux_cnt > FLS_VEC_SZ || m.aux_off + n_vals > nv) {
				return "vector value range outside the chunk's value strings";
			}
			next += n_vals;
			if ((rle && m.width > 16) || (uint64_t)m.data_off16 * 16 + blob > ch.seg_size[FLS_SEG_DATA] || blob > ch.max_blob_bytes) {
				return "vector entry point or width out of range";
			}
			if (!rle) { // exception positions ascend strictly and stay inside the vector
				const uint16_t* pos = reinterpret_cast<const uint16_t*>(data_seg + (uint64_t)m.data_off16 * 16);
				for (uint32_t k = 0; k < m.aux_cnt; ++k) {
					if (pos[k] >= FLS_VEC_SZ || (k && pos[k] <= pos[k - 1])) {
						return "exception positions out of range or not ascending";
					}
				}
			}
		}
		if (ch.encoding == FLS_ENC_RLE_STR && ch.idx_bits != 16) {
			return "RLE_STR idx_bits must be 16";
		}
		return next == nv ? nullptr : "vector value ranges do not cover the chunk's value strings";
	}
	const uint64_t n = dict ? ch.aux_count : (uint64_t)ch.n_vec * FLS_VEC_SZ;
	if (n == 0 || ch.aux_count != n) {
		return "string count disagrees with the chunk descriptor";
	}
	if (ch.seg_size[FLS_SEG_AUX1] != n * 4 * (fsst ? 2 : 1)) {
		return "string offset segment has the wrong size";
	}
	if (dict) {
		if (ch.idx_bits != 8 && ch.idx_bits != 16 && ch.idx_bits != 32) {
			return "dictionary idx_bits must be 8, 16 or 32";
		}
		for (uint32_t v = 0; v < ch.n_vec; ++v) {
			const fls_vec_meta& m = meta[v];
			if (m.width > ch.idx_bits || (uint64_t)m.data_off16 * 16 + 128u * m.width > ch.seg_size[FLS_SEG_DATA] ||
			    128u * m.width > ch.max_blob_bytes || m.aux_cnt != 0) {
				return "code vector entry point or width out of range";
			}
		}
	}
	const uint64_t table_bytes = fsst ? FLS_FSST_TABLE_BYTES(ch.encoding) : 0;
	if (ch.seg_size[FLS_SEG_AUX0] < table_bytes) {
		return "FSST table missing";
	}
	const uint64_t n_bytes = ch.seg_size[FLS_SEG_AUX0] - table_bytes;
	uint32_t prev = 0;
	for (uint64_t i = 0; i < n; ++i) {
		if (aux1[i] < prev) {
			return "string end offsets are not monotone";
		}
		prev = aux1[i];
	}
	if (prev != n_bytes) {
		return "string end offsets do not cover the byte segment";
	}
	if (!fsst) {
		return ch.const_val == n_bytes ? nullptr : "decoded byte count disagrees with the byte segment";
	}
	if (fsst12) {
		// 12-bit codes: every symbol zero padded beyond its length (the device ORs symbols into the output word unmasked),
		// codes < 256 are the bytes themselves, streams decode to the promised lengths
		const fls_fsst12_table* t12 = reinterpret_cast<const fls_fsst12_table*>(aux0);
		for (int c = 0; c < 4096; ++c) {
			const unsigned len = t12->len[c];
			if (len == 0 || len > 8 || (len < 8 && (t12->symbol[c] >> (8 * len)) != 0)) {
				return "FSST12 symbol length out of range or symbol not zero padded";
			}
			if (c < 256 && (len != 1 || t12->symbol[c] != (uint64_t)c)) {
				return "FSST12 codes below 256 must stand for their own byte";
			}
		}
		const uint8_t* codes12 = aux0 + table_bytes;
		const uint32_t* dec_end12 = aux1 + n;
		uint32_t cprev12 = 0, dprev12 = 0;
		for (uint64_t i = 0; i < n; ++i) {
			const uint64_t len = flshost::fsst12_decoded_length(*t12, codes12 + cprev12, aux1[i] - cprev12);
			if (len == ~0ull || dec_end12[i] < dprev12 || dec_end12[i] - dprev12 != len) {
				return "FSST12 decoded end offsets disagree with the code stream";
			}
			cprev12 = aux1[i];
			dprev12 = dec_end12[i];
		}
		return ch.const_val == dprev12 ? nullptr : "decoded byte count disagrees with the decoded end offsets";
	}
	const fls_fsst_table* tab = reinterpret_cast<const fls_fsst_table*>(aux0);
	for (int c = 0; c < 256; ++c) {
		if (tab->len[c] == 0 || tab->len[c] > 8) {
			return "FSST symbol length out of range";
		}
	}
	const uint8_t* codes = aux0 + table_bytes;
	const uint32_t* dec_end = aux1 + n;
	uint32_t cprev = 0, dprev = 0;
	for (uint64_t i = 0; i < n; ++i) {
		uint64_t len = 0;
		for (uint32_t p = cprev; p < aux1[i]; ++p) {
			if (codes[p] == FLS_FSST_ESC) {
				if (++p >= aux1[i]) {
					return "FSST escape without its literal";
				}
				++len;
			} else {
				len += tab->len[codes[p]];
			}
		}
		if (dec_end[i] < dprev || dec_end[i] - dprev != len) {
			return "FSST decoded end offsets disagree with the code stream";
		}
		cprev = aux1[i];
		dprev = dec_end[i];
	}
	return ch.const_val == dprev ? nullptr : "decoded byte count disagrees with the decoded end offsets";
}

// Everything about one chunk that opening a table checks: extents, descriptor consistency, per-vector entry points, exception
// positions, string offsets / FSST streams.  Reads the image only: chunks are validated in parallel (fls_table_parse).
static const char* validate_chunk(const flsgpu_table* t, uint32_t r, uint32_t c) {
	const fls_file_header& h = t->header;
	const fls_rowgroup_desc& rg = t->rowgroups[r];
	const fls_chunk_desc& ch = t->chunk(r, c);
	const fls_column_desc& col = t->columns[c];
	if (ch.offset % FLS_ALIGN || ch.offset > h.footer_offset || ch.size > h.footer_offset - ch.offset) {
		return "extent outside the data region";
	}
	if (ch.encoding >= FLS_ENC_COUNT || ch.phys_bits != col.phys_bits || ch.n_vec != rg.n_vec || ch.n_tuples != rg.n_tuples) {
		return "descriptor disagrees with the row group / column";
	}
	for (int s = 0; s < FLS_SEG_COUNT; ++s) {
		if (ch.seg_off[s] % FLS_ALIGN || ch.seg_off[s] > ch.size || ch.seg_size[s] > ch.size - ch.seg_off[s]) {
			return "segment outside the chunk";
		}
	}
	if (ch.is_equal) { // EXP_EQUAL: a copy of an EARLIER column's descriptor of the same row group, never a chain
		const uint32_t src = ch.equal_col[0] | ((uint32_t)ch.equal_col[1] << 8);
		if (src >= c || col.phys_bits == 128) {
			return "EXP_EQUAL must refer to an earlier numeric column";
		}
		fls_chunk_desc want = t->chunk(r, src);
		if (want.is_equal || t->columns[src].phys_bits != col.phys_bits) {
			return "EXP_EQUAL source is itself a reference or has another physical width";
		}
		want.is_equal = 1;
		want.equal_col[0] = ch.equal_col[0];
		want.equal_col[1] = ch.equal_col[1];
		if (std::memcmp(&want, &ch, sizeof(fls_chunk_desc)) != 0) {
			return "EXP_EQUAL descriptor differs from its source chunk's";
		}
	}
	if (ch.seg_size[FLS_SEG_META] != (uint64_t)ch.n_vec * sizeof(fls_vec_meta)) {
		return "meta segment has the wrong size";
	}
	if (is_string_enc(ch.encoding) != (col.phys_bits == 128)) {
		return "string encoding on a numeric column or vice versa";
	}
	if (col.phys_bits == 128) {
		const char* why = validate_string_chunk(t, ch);
		if (why) {
			return why;
		}
		return nullptr;
	}
	const unsigned T = ch.phys_bits, esz = T / 8;
	const fls_vec_meta* meta = reinterpret_cast<const fls_vec_meta*>(t->bytes + ch.offset + ch.seg_off[FLS_SEG_META]);
	const bool patched = is_patched(ch.encoding);
	unsigned lanes_T = T;
	if (ch.encoding == FLS_ENC_DICT) {
		lanes_T = ch.idx_bits;
		if (lanes_T != 8 && lanes_T != 16 && lanes_T != 32) {
			return "DICT idx_bits must be 8, 16 or 32";
		}
		if (ch.seg_size[FLS_SEG_AUX0] != (uint64_t)ch.aux_count * esz || ch.aux_count == 0) {
			return "dictionary segment has the wrong size";
		}
	} else if (ch.encoding == FLS_ENC_RLE) {
		lanes_T = 16;
		if (ch.seg_size[FLS_SEG_AUX0] != (uint64_t)ch.aux_count * esz) {
			return "run value segment has the wrong size";
		}
	} else if (ch.encoding == FLS_ENC_CROSS_RLE) {
		if (ch.seg_size[FLS_SEG_AUX0] != (uint64_t)ch.aux_count * esz || ch.seg_size[FLS_SEG_AUX1] != (uint64_t)ch.aux_count * 4) {
			return "CROSS_RLE segments have the wrong size";
		}
		const uint32_t* lens = reinterpret_cast<const uint32_t*>(t->bytes + ch.offset + ch.seg_off[FLS_SEG_AUX1]);
		uint64_t sum = 0;
		for (uint32_t i = 0; i < ch.aux_count; ++i) {
			sum += lens[i];
		}
		if (sum != (uint64_t)ch.n_vec * FLS_VEC_SZ) {
			return "CROSS_RLE run lengths do not cover the chunk";
		}
	}
	if ((ch.encoding == FLS_ENC_ALP || ch.encoding == FLS_ENC_ALP_RD) && !(col.is_float && T >= 32)) {
		return "ALP encodings need a FLOAT/DOUBLE column";
	}
	const uint8_t* data_seg = t->bytes + ch.offset + ch.seg_off[FLS_SEG_DATA];
	uint64_t n_exc_seen = 0;
	for (uint32_t v = 0; v < ch.n_vec; ++v) {
		const fls_vec_meta& m = meta[v];
		if (m.width > lanes_T || (ch.encoding == FLS_ENC_ALP_RD && (m.width2 == 0 || m.width2 > 3 || m.width + m.width2 > 64 + 3))) {
			return "vector bit width out of range";
		}
		if ((uint64_t)m.data_off16 * 16 + blob_bytes(ch, m) > ch.seg_size[FLS_SEG_DATA]) {
			return "vector entry point outside the data segment";
		}
		if (blob_bytes(ch, m) > ch.max_blob_bytes) {
			return "vector blob larger than the chunk's max_blob_bytes";
		}
		if (m.n_valid == 0 || m.n_valid > FLS_VEC_SZ) {
			return "vector n_valid out of range";
		}
		if (m.aux_cnt > FLS_VEC_SZ || (!patched && ch.encoding != FLS_ENC_RLE && m.aux_cnt != 0)) {
			return "vector exception/run count out of range";
		}
		if (ch.encoding == FLS_ENC_RLE && (uint64_t)m.aux_off + m.aux_cnt > ch.aux_count) {
			return "vector run range outside the run value segment";
		}
		if (patched && m.aux_cnt) {
			const uint16_t* exc_pos = reinterpret_cast<const uint16_t*>(data_seg + (uint64_t)m.data_off16 * 16 + packed_bytes(ch, m) +
			                                                            FLS_PAD16((uint32_t)m.aux_cnt * esz));
			for (uint32_t k = 0; k < m.aux_cnt; ++k) {
				if (exc_pos[k] >= FLS_VEC_SZ) {
					return "exception position out of range";
				}
			}
			n_exc_seen += m.aux_cnt;
		}
		if (ch.encoding == FLS_ENC_ALP && (m.e > (T == 64 ? 18 : 10) || m.f > m.e)) {
			return "ALP exponent/factor out of range";
		}
		if (ch.encoding == FLS_ENC_RLE && m.aux_cnt == 0) {
			return "RLE vector without runs";
		}
	}
	if (patched && n_exc_seen != ch.aux_count) {
		return "exception count disagrees with the chunk descriptor";
	}
	return nullptr;
}

int fls_table_parse(flsgpu_table* t) {
	char msg[256];
	if (t->size < sizeof(fls_file_header)) {
		return fls_set_error(FLSGPU_ERR_INVALID, "file smaller than its header");
	}
	std::memcpy(&t->header, t->bytes, sizeof(fls_file_header));
	const fls_file_header& h = t->header;
	if (std::memcmp(h.magic, FLS_MAGIC, 8) != 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "bad magic: not an flsgpu container");
	}
	if (h.version != FLS_VERSION) {
		return fls_set_error(FLSGPU_ERR_INVALID, "unsupported container version");
	}
	if (h.n_columns == 0 || h.vec_per_rowgroup == 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "container declares zero columns or zero vectors per row group");
	}
	// every count is bounded by the file size before it enters a product (a hostile header must not wrap the arithmetic)
	const uint64_t max_chunks = t->size / sizeof(fls_chunk_desc);
	if (h.n_columns > max_chunks || h.n_rowgroups > max_chunks || (h.n_rowgroups && h.n_columns > max_chunks / h.n_rowgroups)) {
		return fls_set_error(FLSGPU_ERR_INVALID, "footer out of bounds or of unexpected size");
	}
	const uint64_t need = (uint64_t)h.n_columns * sizeof(fls_column_desc) + (uint64_t)h.n_rowgroups * sizeof(fls_rowgroup_desc) +
	                      (uint64_t)h.n_rowgroups * h.n_columns * sizeof(fls_chunk_desc);
	if (h.footer_offset > t->size || h.footer_size != need || need > t->size - h.footer_offset || (h.footer_offset % 8)) {
		return fls_set_error(FLSGPU_ERR_INVALID, "footer out of bounds or of unexpected size");
	}
	const uint8_t* f = t->bytes + h.footer_offset;
	t->columns = reinterpret_cast<const fls_column_desc*>(f);
	f += (uint64_t)h.n_columns * sizeof(fls_column_desc);
	t->rowgroups = reinterpret_cast<const fls_rowgroup_desc*>(f);
	f += (uint64_t)h.n_rowgroups * sizeof(fls_rowgroup_desc);
	t->chunks = reinterpret_cast<const fls_chunk_desc*>(f);

	for (uint32_t c = 0; c < h.n_columns; ++c) {
		const unsigned pb = t->columns[c].phys_bits;
		if (pb != 8 && pb != 16 && pb != 32 && pb != 64 && pb != 128) {
			std::snprintf(msg, sizeof(msg), "column %u: physical width must be 8, 16, 32, 64 or 128 (strings) bits", c);
			return fls_set_error(FLSGPU_ERR_INVALID, msg);
		}
	}
	uint64_t tuples = 0;
	t->n_vectors = 0;
	for (uint32_t r = 0; r < h.n_rowgroups; ++r) {
		const fls_rowgroup_desc& rg = t->rowgroups[r];
		if (rg.n_tuples == 0 || rg.n_vec != (rg.n_tuples + FLS_VEC_SZ - 1) / FLS_VEC_SZ || rg.n_vec > h.vec_per_rowgroup ||
		    rg.first_tuple != tuples) {
			std::snprintf(msg, sizeof(msg), "row group %u: inconsistent tuple/vector counts", r);
			return fls_set_error(FLSGPU_ERR_INVALID, msg);
		}
		if (r + 1 < h.n_rowgroups && rg.n_vec != h.vec_per_rowgroup) {
			std::snprintf(msg, sizeof(msg), "row group %u: only the last row group may be short", r);
			return fls_set_error(FLSGPU_ERR_INVALID, msg);
		}
		tuples += rg.n_tuples;
		t->n_vectors += rg.n_vec;
	}
	// The per-chunk checks walk every entry point, exception list and FSST code stream - O(compressed bytes): 0.29 s for a
	// 196 MB SF1 lineitem on one core.  Row groups are independent, so they are spread over the host's cores; the first
	// failure in (row group, column) order is the one reported.
	{
		const unsigned n_threads = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>({(uint64_t)std::thread::hardware_concurrency(), 32u,
		                                                                                 (uint64_t)h.n_rowgroups, h.footer_offset >> 22}));
		std::atomic<uint32_t> next {0};
		std::atomic<uint64_t> first_bad {~0ull};
		std::vector<const char*> why(n_threads, nullptr);
		std::vector<uint64_t> where(n_threads, ~0ull);
		auto work = [&](unsigned k) {
			for (;;) {
				const uint32_t r = next.fetch_add(1, std::memory_order_relaxed);
				if (r >= h.n_rowgroups || (uint64_t)r * h.n_columns > first_bad.load(std::memory_order_relaxed)) {
					return;
				}
				for (uint32_t c = 0; c < h.n_columns; ++c) {
					const char* w = validate_chunk(t, r, c);
					if (w) {
						const uint64_t at = (uint64_t)r * h.n_columns + c;
						if (at < where[k]) {
							where[k] = at;
							why[k] = w;
						}
						uint64_t cur = first_bad.load(std::memory_order_relaxed);
						while (at < cur && !first_bad.compare_exchange_weak(cur, at)) {
						}
						break;
					}
				}
			}
		};
		const auto t_dbg0 = std::chrono::steady_clock::now();
		std::vector<std::thread> pool;
		for (unsigned k = 1; k < n_threads; ++k) {
			pool.emplace_back(work, k);
		}
		work(0);
		for (auto& th : pool) {
			th.join();
		}
		if (std::getenv("FLSGPU_TRACE_OPEN")) {
			std::fprintf(stderr, "flsgpu: validated %u row groups on %u threads in %.3f s\n", h.n_rowgroups, n_threads,
			             std::chrono::duration<double>(std::chrono::steady_clock::now() - t_dbg0).count());
		}
		const uint64_t bad = first_bad.load();
		if (bad != ~0ull) {
			for (unsigned k = 0; k < n_threads; ++k) {
				if (where[k] == bad) {
					std::snprintf(msg, sizeof(msg), "chunk (rg %u, col %u): %s", (unsigned)(bad / h.n_columns), (unsigned)(bad % h.n_columns), why[k]);
					return fls_set_error(FLSGPU_ERR_INVALID, msg);
				}
			}
		}
	}
	if (tuples != h.n_tuples) {
		return fls_set_error(FLSGPU_ERR_INVALID, "row group tuple counts do not add up to the table's");
	}
	return FLSGPU_OK;
}

extern "C" {

const char* flsgpu_last_error(void) {
	return g_last_error.c_str();
}
const char* flsgpu_version(void) {
	return "flsgpu 0.1 (sm_100a)";
}

int flsgpu_table_open_memory(const void* bytes, uint64_t size, int copy, flsgpu_table** out) {
	if (!bytes || !out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_open_memory: null argument");
	}
	auto t = std::make_unique<flsgpu_table>();
	if (copy) {
		t->owned = std::make_shared<FlsImage>(static_cast<const uint8_t*>(bytes), static_cast<const uint8_t*>(bytes) + size);
		t->bytes = t->owned->data();
	} else {
		t->bytes = static_cast<const uint8_t*>(bytes);
	}
	t->size = size;
	const int rc = fls_table_parse(t.get());
	if (rc) {
		return rc;
	}
	*out = t.release();
	return FLSGPU_OK;
}

int flsgpu_table_open(const char* path, flsgpu_table** out) {
	if (!path || !out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_open: null argument");
	}
	FILE* fp = std::fopen(path, "rb");
	if (!fp) {
		return fls_set_error(FLSGPU_ERR_IO, (std::string("cannot open file: ") + path).c_str());
	}
	std::fseek(fp, 0, SEEK_END);
	const long sz = std::ftell(fp);
	std::fseek(fp, 0, SEEK_SET);
	auto t = std::make_unique<flsgpu_table>();
	t->owned = std::make_shared<FlsImage>(sz > 0 ? (size_t)sz : 0);
	// the image is read by several threads at once (pread of disjoint pieces): out of the page cache a single reader is bound
	// by its own page faults and copies, 0.13 s for a 196 MB file
	size_t n = 0;
	if (sz > 0) {
		const int fd = fileno(fp);
		const size_t total = (size_t)sz;
		const unsigned n_readers = (unsigned)std::max<size_t>(1, std::min<size_t>({(size_t)std::thread::hardware_concurrency(), 16u, total >> 24}));
		const size_t piece = (total + n_readers - 1) / n_readers;
		std::vector<size_t> got(n_readers, 0);
		auto read_piece = [&](unsigned k) {
			const size_t begin = std::min(total, (size_t)k * piece), end = std::min(total, begin + piece);
			size_t at = begin;
			while (at < end) {
				const ssize_t r = pread(fd, t->owned->data() + at, end - at, (off_t)at);
				if (r <= 0) {
					break;
				}
				at += (size_t)r;
			}
			got[k] = at - begin;
		};
		std::vector<std::thread> readers;
		for (unsigned k = 1; k < n_readers; ++k) {
			readers.emplace_back(read_piece, k);
		}
		read_piece(0);
		for (auto& th : readers) {
			th.join();
		}
		for (size_t g : got) {
			n += g;
		}
	}
	std::fclose(fp);
	if (sz < 0 || n != (size_t)sz) {
		return fls_set_error(FLSGPU_ERR_IO, (std::string("short read: ") + path).c_str());
	}
	// footer written beside the data (COPY ... (FORMAT fls, inline_footer false), reference fls_writer.cpp:229-230): append it
	if (t->owned->size() >= sizeof(fls_file_header)) {
		fls_file_header h;
		std::memcpy(&h, t->owned->data(), sizeof(h));
		if (std::memcmp(h.magic, FLS_MAGIC, 8) == 0 && h.reserved[0] == 1 && h.footer_offset == t->owned->size() &&
		    h.footer_size <= (1ull << 40)) {
			const std::string fpath = std::string(path) + ".footer";
			FILE* ff = std::fopen(fpath.c_str(), "rb");
			if (!ff) {
				return fls_set_error(FLSGPU_ERR_IO, (std::string("cannot open footer file: ") + fpath).c_str());
			}
			const size_t at = t->owned->size();
			t->owned->resize(at + (size_t)h.footer_size);
			const size_t k = std::fread(t->owned->data() + at, 1, (size_t)h.footer_size, ff);
			std::fclose(ff);
			if (k != (size_t)h.footer_size) {
				return fls_set_error(FLSGPU_ERR_IO, (std::string("short read: ") + fpath).c_str());
			}
		}
	}
	t->bytes = t->owned->data();
	t->size = t->owned->size();
	const int rc = fls_table_parse(t.get());
	if (rc) {
		return rc;
	}
	*out = t.release();
	return FLSGPU_OK;
}

// Another handle over the same host image (no copy, no second validation pass): one handle per device of a single-process
// multi-GPU scan.  The image lives as long as any handle that owns it does.
int flsgpu_table_share(const flsgpu_table* src, flsgpu_table** out) {
	if (!src || !out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_share: null argument");
	}
	auto t = std::make_unique<flsgpu_table>();
	t->bytes = src->bytes;
	t->size = src->size;
	t->owned = src->owned;
	t->header = src->header;
	t->columns = src->columns;
	t->rowgroups = src->rowgroups;
	t->chunks = src->chunks;
	t->n_vectors = src->n_vectors;
	t->device_index = src->device_index;
	*out = t.release();
	return FLSGPU_OK;
}

void flsgpu_table_close(flsgpu_table* t) {
	if (t) {
		fls_device_release(t);
		delete t;
	}
}

int flsgpu_table_info_get(const flsgpu_table* t, flsgpu_table_info* out) {
	if (!t || !out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_info_get: null argument");
	}
	out->n_columns = t->header.n_columns;
	out->n_rowgroups = t->header.n_rowgroups;
	out->vec_per_rowgroup = t->header.vec_per_rowgroup;
	out->reserved = 0;
	out->n_tuples = t->header.n_tuples;
	out->n_vectors = t->n_vectors;
	out->file_bytes = t->size;
	return FLSGPU_OK;
}

// Statistics of a string column for the host database's optimizer, where they come cheap: chunks whose distinct strings sit
// uncompressed in AUX0 (DICT_STR dictionaries, STR_CONSTANT, the value strings of RLE_STR / CROSS_RLE_STR / FREQUENCY_STR) are
// summarised from those few strings; any other chunk (one string per row, or an FSST-compressed dictionary) makes the
// statistics unavailable rather than costing a decode of the column.
int flsgpu_column_string_stats(const flsgpu_table* t, uint32_t col, flsgpu_string_stats* out) {
	if (!t || !out || col >= t->header.n_columns || t->columns[col].phys_bits != 128) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_column_string_stats: null argument or not a string column");
	}
	flsgpu_string_stats st {};
	const uint8_t *min_p = nullptr, *max_p = nullptr;
	uint32_t min_n = 0, max_n = 0;
	auto less = [](const uint8_t* a, uint32_t an, const uint8_t* b, uint32_t bn) { // bytewise, the shorter first on a tie
		const int c = std::memcmp(a, b, std::min(an, bn));
		return c < 0 || (c == 0 && an < bn);
	};
	auto take = [&](const uint8_t* p, uint32_t n) {
		st.max_length = std::max(st.max_length, n);
		for (uint32_t i = 0; i < n && !st.has_non_ascii; ++i) {
			st.has_non_ascii = p[i] >= 0x80u;
		}
		if (!min_p || less(p, n, min_p, min_n)) {
			min_p = p;
			min_n = n;
		}
		if (!max_p || less(max_p, max_n, p, n)) {
			max_p = p;
			max_n = n;
		}
	};
	for (uint32_t r = 0; r < t->header.n_rowgroups; ++r) {
		const fls_chunk_desc& ch = t->chunk(r, col);
		const uint8_t* aux0 = t->bytes + ch.offset + ch.seg_off[FLS_SEG_AUX0];
		const uint32_t* ends = reinterpret_cast<const uint32_t*>(t->bytes + ch.offset + ch.seg_off[FLS_SEG_AUX1]);
		if (ch.encoding == FLS_ENC_STR_CONSTANT) {
			take(aux0, ch.aux_count);
		} else if (ch.encoding == FLS_ENC_DICT_STR || FLS_ENC_IS_STR_VALUES(ch.encoding)) {
			for (uint32_t k = 0; k < ch.aux_count; ++k) {
				const uint32_t start = k ? ends[k - 1] : 0u;
				take(aux0 + start, ends[k] - start);
			}
		} else {
			*out = flsgpu_string_stats {};
			return FLSGPU_OK; // available = 0
		}
	}
	if (min_p) {
		st.available = 1;
		st.min_length = std::min(min_n, 8u);
		st.max_value_length = std::min(max_n, 8u);
		std::memcpy(st.min_value, min_p, st.min_length);
		std::memcpy(st.max_value, max_p, st.max_value_length);
	}
	*out = st;
	return FLSGPU_OK;
}

// Vector forms (fls_format.h FLS_VEC_FORM_*): read off the entry points, no decode.
int flsgpu_column_vector_forms(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, flsgpu_vector_form* out,
                               uint64_t* n_flat) {
	if (!t || !out || col >= t->header.n_columns || n_rg == 0 || rg_begin >= t->header.n_rowgroups ||
	    n_rg > t->header.n_rowgroups - rg_begin) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_column_vector_forms: null argument, column or row-group range out of bounds");
	}
	const fls_column_desc& cd = t->columns[col];
	const unsigned T = cd.phys_bits;
	const uint32_t vpr = t->header.vec_per_rowgroup;
	uint32_t flat = 0;
	auto extend = [&](uint64_t word) -> int64_t { // physical word -> 64 bits by the column's signedness
		if (T >= 64 || cd.is_float) {
			return (int64_t)word;
		}
		const uint64_t m = (1ull << T) - 1;
		word &= m;
		return (cd.is_signed && (word >> (T - 1))) ? (int64_t)(word | ~m) : (int64_t)word;
	};
	for (uint32_t r = 0; r < n_rg; ++r) {
		const fls_chunk_desc& ch = t->chunk(rg_begin + r, col);
		const fls_vec_meta* meta = reinterpret_cast<const fls_vec_meta*>(t->bytes + ch.offset + ch.seg_off[FLS_SEG_META]);
		for (uint32_t v = 0; v < vpr; ++v) {
			flsgpu_vector_form& f = out[(uint64_t)r * vpr + v];
			f.start = f.increment = 0;
			f.form = FLS_VEC_FORM_CONSTANT;
			f.flat_slot = 0xffffffffu;
			if (v >= ch.n_vec) {
				continue;
			}
			if (T != 128 && ch.encoding == FLS_ENC_CONSTANT) {
				f.start = extend(ch.const_val);
				continue;
			}
			const fls_vec_meta& m = meta[v];
			if (T != 128 && ch.encoding == FLS_ENC_FFOR && m.width == 0 && m.aux_cnt == 0) {
				f.start = extend(m.base);
				continue;
			}
			if (T != 128 && !cd.is_float && ch.encoding == FLS_ENC_DELTA && m.width == 0 && m.e == FLS_VEC_FORM_SEQUENCE) {
				// increment: the delta as a signed T-bit number (a decreasing unsigned column steps by -k, not by 2^T - k);
				// the progression must stay inside the column's type and inside int64 (what DuckDB's Vector::Sequence takes)
				const uint64_t mask = T >= 64 ? ~0ull : (1ull << T) - 1;
				const uint64_t dw = m.base & mask;
				const int64_t inc = (T < 64 && (dw >> (T - 1))) ? (int64_t)(dw | ~mask) : (int64_t)dw;
				// last = start + 1023 * inc, all in 64 bits with overflow detection (an unsigned column's values are kept as
				// int64 too, so anything beyond INT64_MAX is out: DuckDB's Vector::Sequence takes int64)
				const uint64_t sw = m.reserved & mask;
				int64_t start, span, last;
				bool ok = cd.is_signed || sw <= (uint64_t)INT64_MAX;
				start = cd.is_signed ? extend(m.reserved) : (int64_t)sw;
				ok = ok && !__builtin_mul_overflow(inc, (int64_t)1023, &span) && !__builtin_add_overflow(start, span, &last);
				if (ok && T < 64) {
					const int64_t lo = cd.is_signed ? -((int64_t)1 << (T - 1)) : 0;
					const int64_t hi = cd.is_signed ? ((int64_t)1 << (T - 1)) - 1 : (int64_t)mask;
					ok = start >= lo && start <= hi && last >= lo && last <= hi;
				} else if (ok && !cd.is_signed) {
					ok = last >= 0;
				}
				if (ok) {
					f.start = start;
					f.increment = inc;
					f.form = inc == 0 ? FLS_VEC_FORM_CONSTANT : FLS_VEC_FORM_SEQUENCE;
					continue;
				}
			}
			f.form = FLS_VEC_FORM_FLAT;
			f.flat_slot = flat++;
		}
	}
	if (n_flat) {
		*n_flat = flat;
	}
	return FLSGPU_OK;
}

int flsgpu_table_column(const flsgpu_table* t, uint32_t col, fls_column_desc* out) {
	if (!t || !out || col >= t->header.n_columns) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_column: column index out of range");
	}
	*out = t->columns[col];
	return FLSGPU_OK;
}

int flsgpu_table_rowgroup(const flsgpu_table* t, uint32_t rg, fls_rowgroup_desc* out) {
	if (!t || !out || rg >= t->header.n_rowgroups) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_rowgroup: row group index out of range");
	}
	*out = t->rowgroups[rg];
	return FLSGPU_OK;
}

int flsgpu_table_chunk(const flsgpu_table* t, uint32_t rg, uint32_t col, fls_chunk_desc* out) {
	if (!t || !out || rg >= t->header.n_rowgroups || col >= t->header.n_columns) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_chunk: index out of range");
	}
	*out = t->chunk(rg, col);
	return FLSGPU_OK;
}

int flsgpu_string_heap_bytes(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* out) {
	if (!t || !out || col >= t->header.n_columns || rg_begin > t->header.n_rowgroups || n_rg > t->header.n_rowgroups - rg_begin) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_string_heap_bytes: column or row-group range out of bounds");
	}
	if (t->columns[col].phys_bits != 128) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_string_heap_bytes: not a string column");
	}
	uint64_t total = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		total += (t->chunk(r, col).const_val + 15u) & ~15ull; // every chunk's slice of the heap starts 16-byte aligned
	}
	*out = total;
	return FLSGPU_OK;
}

int flsgpu_string_dict_keys(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* out) {
	if (!t || !out || col >= t->header.n_columns || rg_begin > t->header.n_rowgroups || n_rg > t->header.n_rowgroups - rg_begin) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_string_dict_keys: column or row-group range out of bounds");
	}
	uint64_t total = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		const fls_chunk_desc& ch = t->chunk(r, col);
		if (!FLS_ENC_IS_STR_DICT(ch.encoding)) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_string_dict_keys: chunk is not dictionary encoded");
		}
		total += ch.aux_count;
	}
	*out = total;
	return FLSGPU_OK;
}

} // extern "C"
