// fls_strings.cpp — string column chunks (VARCHAR / BLOB): STR_CONSTANT, STR_UNCOMPRESSED, DICT_STR, FSST, FSST_DICT, FSST12,
// FSST12_DICT, RLE_STR, CROSS_RLE_STR, FREQUENCY_STR.
// Layouts: include/fls_format.h ("String columns").  Stands in for the string side of fastlanes::RowGroupWriter
// (reference src/writer/fls_writer.cpp:14-39; fls_view_writer.cpp:61-94 hands VARCHAR/BLOB to the library as
// offsets + bytes); which scheme a chunk gets is encoder policy (absent library), here: smallest estimated size.
#include "fls_encoder_internal.hpp"
#include "fls_fsst.hpp"

#include <algorithm>
#include <cstring>
#include <memory>
#include <string_view>
#include <unordered_map>

namespace flshost {

namespace {

struct StrChunk {
	const uint64_t* offsets; // n_tuples + 1, absolute positions in bytes
	const uint8_t* bytes;
	uint32_t n_tuples, n_vec;
	std::string_view at(uint32_t i) const {
		return std::string_view(reinterpret_cast<const char*>(bytes + offsets[i]), (size_t)(offsets[i + 1] - offsets[i]));
	}
	uint16_t valid_in(uint32_t v) const {
		const uint32_t left = n_tuples - v * kVec;
		return (uint16_t)(left < (uint32_t)kVec ? left : (uint32_t)kVec);
	}
	fls_chunk_desc base_desc(int enc) const {
		fls_chunk_desc d {};
		d.encoding = (uint8_t)enc;
		d.phys_bits = 128;
		d.n_vec = n_vec;
		d.n_tuples = n_tuples;
		return d;
	}
};

void plain_meta(const StrChunk& c, ChunkBuilder& b) { // entry points of encodings without per-vector blobs
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		b.meta.push_back(m);
	}
}

void append_u32(std::vector<uint8_t>& dst, const std::vector<uint32_t>& src) {
	const size_t o = dst.size();
	dst.resize(o + src.size() * 4);
	if (!src.empty()) {
		std::memcpy(dst.data() + o, src.data(), src.size() * 4);
	}
}

void encode_str_constant(const StrChunk& c, EncodedChunk& out) {
	ChunkBuilder b;
	plain_meta(c, b);
	const std::string_view s = c.at(0);
	b.aux0.assign(s.begin(), s.end());
	fls_chunk_desc d = c.base_desc(FLS_ENC_STR_CONSTANT);
	d.aux_count = (uint32_t)s.size();
	d.const_val = s.size();
	b.finish(out, d);
}

void encode_str_uncompressed(const StrChunk& c, EncodedChunk& out) {
	ChunkBuilder b;
	plain_meta(c, b);
	const uint64_t total = c.offsets[c.n_tuples] - c.offsets[0];
	b.aux0.assign(c.bytes + c.offsets[0], c.bytes + c.offsets[c.n_tuples]);
	std::vector<uint32_t> ends((size_t)c.n_vec * kVec, (uint32_t)total); // padding rows: empty strings at the end
	for (uint32_t i = 0; i < c.n_tuples; ++i) {
		ends[i] = (uint32_t)(c.offsets[i + 1] - c.offsets[0]);
	}
	append_u32(b.aux1, ends);
	fls_chunk_desc d = c.base_desc(FLS_ENC_STR_UNCOMPRESSED);
	d.aux_count = (uint32_t)ends.size();
	d.const_val = total;
	b.finish(out, d);
}

// AUX0 = table | code streams, AUX1 = comp_end[n] | dec_end[n] for the n strings given by (offsets, bytes)
template <class Encoder, class Table>
void fsst_segments_t(const uint64_t* offsets, const uint8_t* bytes, uint32_t n, uint32_t n_padded, ChunkBuilder& b, uint64_t* decoded_total) {
	auto enc_holder = std::make_unique<Encoder>(); // the FSST12 encoder carries a 36 KiB table: keep it off the stack
	Encoder& enc = *enc_holder;
	enc.train(offsets, bytes, n);
	b.aux0.resize(sizeof(Table));
	std::memcpy(b.aux0.data(), &enc.table, sizeof(Table));
	std::vector<uint8_t> codes;
	codes.reserve((size_t)(offsets[n] - offsets[0]) / 2 + 16);
	std::vector<uint32_t> comp_end(n_padded), dec_end(n_padded);
	for (uint32_t i = 0; i < n; ++i) {
		enc.compress(bytes + offsets[i], offsets[i + 1] - offsets[i], codes);
		comp_end[i] = (uint32_t)codes.size();
		dec_end[i] = (uint32_t)(offsets[i + 1] - offsets[0]);
	}
	for (uint32_t i = n; i < n_padded; ++i) {
		comp_end[i] = (uint32_t)codes.size();
		dec_end[i] = (uint32_t)(offsets[n] - offsets[0]);
	}
	b.aux0.insert(b.aux0.end(), codes.begin(), codes.end());
	append_u32(b.aux1, comp_end);
	append_u32(b.aux1, dec_end);
	*decoded_total = offsets[n] - offsets[0];
}

void fsst_segments(bool twelve, const uint64_t* offsets, const uint8_t* bytes, uint32_t n, uint32_t n_padded, ChunkBuilder& b,
                   uint64_t* decoded_total) {
	if (twelve) {
		fsst_segments_t<Fsst12Encoder, fls_fsst12_table>(offsets, bytes, n, n_padded, b, decoded_total);
	} else {
		fsst_segments_t<FsstEncoder, fls_fsst_table>(offsets, bytes, n, n_padded, b, decoded_total);
	}
}

void encode_str_fsst(const StrChunk& c, bool twelve, EncodedChunk& out) {
	ChunkBuilder b;
	plain_meta(c, b);
	uint64_t total = 0;
	fsst_segments(twelve, c.offsets, c.bytes, c.n_tuples, c.n_vec * kVec, b, &total);
	fls_chunk_desc d = c.base_desc(twelve ? FLS_ENC_FSST12 : FLS_ENC_FSST);
	d.aux_count = c.n_vec * kVec;
	d.const_val = total;
	b.finish(out, d);
}

// codes of one vector, FFOR-packed in idx_bits lanes (same blob as numeric DICT: dec_dict_opr<KEY, IDX>::Index())
void pack_codes(const uint32_t* codes, unsigned idx_bits, ChunkBuilder& b, fls_vec_meta& m) {
	uint32_t cmin = ~0u, cmax = 0;
	for (int i = 0; i < kVec; ++i) {
		cmin = std::min(cmin, codes[i]);
		cmax = std::max(cmax, codes[i]);
	}
	const unsigned W = bits_needed<uint32_t>(cmax - cmin);
	m.data_off16 = (uint32_t)(b.data.size() / 16);
	m.base = cmin;
	m.width = (uint8_t)W;
	const size_t old = b.data.size();
	b.data.resize(old + (size_t)W * 128);
	if (!W) {
		return;
	}
	if (idx_bits == 8) {
		uint8_t r[kVec];
		for (int i = 0; i < kVec; ++i) r[i] = (uint8_t)(codes[i] - cmin);
		pack_vector<uint8_t>(r, W, b.data.data() + old);
	} else if (idx_bits == 16) {
		uint16_t r[kVec];
		for (int i = 0; i < kVec; ++i) r[i] = (uint16_t)(codes[i] - cmin);
		pack_vector<uint16_t>(r, W, reinterpret_cast<uint16_t*>(b.data.data() + old));
	} else {
		uint32_t r[kVec];
		for (int i = 0; i < kVec; ++i) r[i] = codes[i] - cmin;
		pack_vector<uint32_t>(r, W, reinterpret_cast<uint32_t*>(b.data.data() + old));
	}
}

struct StrDict {
	std::vector<uint32_t> codes;     // per tuple
	std::vector<uint64_t> key_off;   // n_keys + 1 offsets into key_bytes
	std::vector<uint8_t> key_bytes;  // keys back to back, in order of first appearance
};

// false when the chunk has more than max_keys distinct strings
bool build_dict(const StrChunk& c, size_t max_keys, StrDict& d) {
	std::unordered_map<std::string_view, uint32_t> map;
	map.reserve(1024);
	d.codes.resize(c.n_tuples);
	d.key_off.assign(1, 0);
	for (uint32_t i = 0; i < c.n_tuples; ++i) {
		const std::string_view s = c.at(i);
		auto it = map.find(s);
		if (it == map.end()) {
			if (map.size() >= max_keys) {
				return false;
			}
			it = map.emplace(s, (uint32_t)map.size()).first;
			d.key_bytes.insert(d.key_bytes.end(), s.begin(), s.end());
			d.key_off.push_back(d.key_bytes.size());
		}
		d.codes[i] = it->second;
	}
	return true;
}

void encode_str_dict(const StrChunk& c, const StrDict& dict, int enc_kind, EncodedChunk& out) {
	const bool fsst = enc_kind != FLS_ENC_DICT_STR;
	const uint32_t n_keys = (uint32_t)dict.key_off.size() - 1;
	const unsigned idx_bits = n_keys <= 256 ? 8 : n_keys <= 65536 ? 16 : 32;
	ChunkBuilder b;
	uint32_t codes[kVec];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const uint32_t n_valid = c.valid_in(v);
		for (uint32_t i = 0; i < (uint32_t)kVec; ++i) { // padding rows repeat the last real code
			codes[i] = dict.codes[(size_t)v * kVec + std::min(i, n_valid - 1)];
		}
		fls_vec_meta m {};
		m.n_valid = (uint16_t)n_valid;
		pack_codes(codes, idx_bits, b, m);
		b.meta.push_back(m);
	}
	fls_chunk_desc d = c.base_desc(enc_kind);
	d.idx_bits = (uint8_t)idx_bits;
	d.aux_count = n_keys;
	if (fsst) {
		uint64_t total = 0;
		fsst_segments(enc_kind == FLS_ENC_FSST12_DICT, dict.key_off.data(), dict.key_bytes.data(), n_keys, n_keys, b, &total);
		d.const_val = total;
	} else {
		b.aux0 = dict.key_bytes;
		std::vector<uint32_t> ends(n_keys);
		for (uint32_t k = 0; k < n_keys; ++k) {
			ends[k] = (uint32_t)dict.key_off[k + 1];
		}
		append_u32(b.aux1, ends);
		d.const_val = dict.key_bytes.size();
	}
	b.finish(out, d);
}

// value strings of the VALUE encodings: bytes back to back + end offsets
struct StrValues {
	std::vector<uint8_t> bytes;
	std::vector<uint32_t> ends;
	void push(std::string_view s) {
		bytes.insert(bytes.end(), s.begin(), s.end());
		ends.push_back((uint32_t)bytes.size());
	}
	void into(ChunkBuilder& b, fls_chunk_desc& d) const {
		b.aux0 = bytes;
		append_u32(b.aux1, ends);
		d.aux_count = (uint32_t)ends.size();
		d.const_val = bytes.size();
	}
};

// row i of the chunk, padding rows of the last vector repeat the last real string
std::string_view padded_at(const StrChunk& c, uint32_t i) {
	return c.at(std::min(i, c.n_tuples - 1));
}

// RLE_STR (dec_rle_map_opr<fls_string_t, u16>, rle_map_kernel.hpp:26-61): per vector its run strings and the run index of every
// row, DELTA-coded like the numeric RLE index
void encode_str_rle(const StrChunk& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_RLE_STR);
	d.idx_bits = 16;
	StrValues vals;
	unsigned maxw = 0;
	uint16_t idxs[kVec];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.aux_off = (uint32_t)vals.ends.size();
		uint16_t run = 0;
		std::string_view prev = padded_at(c, v * kVec);
		vals.push(prev);
		idxs[0] = 0;
		for (int i = 1; i < kVec; ++i) {
			const std::string_view cur = padded_at(c, v * kVec + i);
			if (cur != prev) {
				++run;
				vals.push(cur);
				prev = cur;
			}
			idxs[i] = run;
		}
		m.aux_cnt = (uint16_t)(run + 1u);
		delta_encode_vector<uint16_t>(idxs, b, m);
		maxw = std::max<unsigned>(maxw, m.width);
		b.meta.push_back(m);
	}
	d.max_width = (uint8_t)maxw;
	vals.into(b, d);
	b.finish(out, d);
}

// CROSS_RLE_STR (dec_cross_rle_opr<fls_string_t>, cross_rle_kernel.hpp:55-97): runs over the whole chunk
void encode_str_cross_rle(const StrChunk& c, EncodedChunk& out) {
	ChunkBuilder b;
	plain_meta(c, b);
	fls_chunk_desc d = c.base_desc(FLS_ENC_CROSS_RLE_STR);
	StrValues vals;
	std::vector<uint32_t> lens;
	const uint32_t total = c.n_vec * kVec;
	for (uint32_t i = 0; i < total;) {
		const std::string_view s = padded_at(c, i);
		uint32_t j = i + 1;
		while (j < total && padded_at(c, j) == s) {
			++j;
		}
		vals.push(s);
		lens.push_back(j - i);
		i = j;
	}
	vals.into(b, d);
	append_u32(b.aux1, lens);
	b.finish(out, d);
}

// FREQUENCY_STR (dec_frequency_str_opr, frequency_kernel.hpp:23-69): per vector the most frequent string once, and the rows that
// differ as (position, string) exceptions in position order
void encode_str_frequency(const StrChunk& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_FREQUENCY_STR);
	StrValues vals;
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		std::unordered_map<std::string_view, uint32_t> count;
		std::string_view freq = padded_at(c, v * kVec);
		uint32_t best = 0;
		for (int i = 0; i < kVec; ++i) {
			const std::string_view s = padded_at(c, v * kVec + i);
			const uint32_t n = ++count[s];
			if (n > best) { // first string to reach the highest count: deterministic
				best = n;
				freq = s;
			}
		}
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.aux_off = (uint32_t)vals.ends.size();
		m.data_off16 = (uint32_t)(b.data.size() / 16);
		vals.push(freq);
		std::vector<uint16_t> pos;
		for (int i = 0; i < kVec; ++i) {
			const std::string_view s = padded_at(c, v * kVec + i);
			if (s != freq) {
				pos.push_back((uint16_t)i);
				vals.push(s);
			}
		}
		m.aux_cnt = (uint16_t)pos.size();
		const size_t o = b.data.size();
		b.data.resize(o + FLS_PAD16(pos.size() * 2), 0);
		if (!pos.empty()) {
			std::memcpy(b.data.data() + o, pos.data(), pos.size() * 2);
		}
		b.meta.push_back(m);
	}
	vals.into(b, d);
	b.finish(out, d);
}

} // namespace

bool encode_string_chunk(const uint64_t* offsets, const uint8_t* bytes, uint32_t n_tuples, int force_encoding, EncodedChunk& out,
                         const char** err) {
	static const char* none = "";
	if (!err) {
		err = &none;
	}
	if (n_tuples == 0) {
		*err = "empty column chunk";
		return false;
	}
	if (offsets[n_tuples] - offsets[0] >= (1ull << 32)) {
		*err = "more than 4 GiB of string bytes in one row group";
		return false;
	}
	StrChunk c {offsets, bytes, n_tuples, (n_tuples + kVec - 1) / kVec};
	int enc = force_encoding;
	StrDict dict;
	bool have_dict = false;
	if (enc < 0) {
		bool all_equal = true;
		const std::string_view first = c.at(0);
		for (uint32_t i = 1; i < n_tuples && all_equal; ++i) {
			all_equal = c.at(i) == first;
		}
		if (all_equal) {
			enc = FLS_ENC_STR_CONSTANT;
		} else {
			// dictionary when at most a quarter of the rows are distinct (codes + keys beat FSST's ~2x on typical text)
			have_dict = build_dict(c, std::max<size_t>(2, n_tuples / 4), dict);
			if (have_dict) {
				enc = dict.key_bytes.size() >= (64u << 10) ? FLS_ENC_FSST_DICT : FLS_ENC_DICT_STR;
			} else {
				enc = (offsets[n_tuples] - offsets[0]) >= 4096 ? FLS_ENC_FSST : FLS_ENC_STR_UNCOMPRESSED;
			}
		}
	}
	switch (enc) {
	case FLS_ENC_STR_CONSTANT:
		for (uint32_t i = 1; i < n_tuples; ++i) {
			if (c.at(i) != c.at(0)) {
				*err = "STR_CONSTANT forced on a non-constant chunk";
				return false;
			}
		}
		encode_str_constant(c, out);
		return true;
	case FLS_ENC_STR_UNCOMPRESSED: encode_str_uncompressed(c, out); return true;
	case FLS_ENC_FSST: encode_str_fsst(c, false, out); return true;
	case FLS_ENC_FSST12: encode_str_fsst(c, true, out); return true;
	case FLS_ENC_DICT_STR:
	case FLS_ENC_FSST_DICT:
	case FLS_ENC_FSST12_DICT:
		if (!have_dict && !build_dict(c, ~(size_t)0 >> 1, dict)) {
			*err = "dictionary too large";
			return false;
		}
		encode_str_dict(c, dict, enc, out);
		return true;
	case FLS_ENC_RLE_STR: encode_str_rle(c, out); return true;
	case FLS_ENC_CROSS_RLE_STR: encode_str_cross_rle(c, out); return true;
	case FLS_ENC_FREQUENCY_STR: encode_str_frequency(c, out); return true;
	default: *err = "encoding not applicable to a string column"; return false;
	}
}

} // namespace flshost
