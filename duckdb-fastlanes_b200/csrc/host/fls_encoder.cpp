// fls_encoder.cpp — integer encodings + encoding selection of the product writer (host C++).
// See fls_encoder.hpp.  Float encodings (ALP / ALP-RD) are in fls_alp.cpp.
#include "fls_encoder.hpp"
#include "fls_encoder_internal.hpp"

#include <algorithm>
#include <cstring>

namespace flshost {

ColumnType column_type_of(int data_type, int decimal_precision) {
	ColumnType t;
	t.data_type = (uint8_t)data_type;
	switch (data_type) {
	case FLS_TYPE_INT8: t.phys_bits = 8; t.is_signed = true; break;
	case FLS_TYPE_INT16: t.phys_bits = 16; t.is_signed = true; break;
	case FLS_TYPE_INT32: t.phys_bits = 32; t.is_signed = true; break;
	case FLS_TYPE_INT64: t.phys_bits = 64; t.is_signed = true; break;
	case FLS_TYPE_UINT8: t.phys_bits = 8; break;
	case FLS_TYPE_UINT16: t.phys_bits = 16; break;
	case FLS_TYPE_UINT32: t.phys_bits = 32; break;
	case FLS_TYPE_UINT64: t.phys_bits = 64; break;
	case FLS_TYPE_FLOAT: t.phys_bits = 32; t.is_float = true; break;
	case FLS_TYPE_DOUBLE: t.phys_bits = 64; t.is_float = true; break;
	case FLS_TYPE_BOOLEAN: t.phys_bits = 8; break;
	case FLS_TYPE_DATE: t.phys_bits = 32; t.is_signed = true; break;
	case FLS_TYPE_TIMESTAMP: t.phys_bits = 64; t.is_signed = true; break;
	case FLS_TYPE_DECIMAL:
		// DuckDB physical width by precision (reference src/writer/translation_utils.cpp:46-58)
		t.phys_bits = decimal_precision <= 4 ? 16 : decimal_precision <= 9 ? 32 : 64;
		t.is_signed = true;
		break;
	case FLS_TYPE_FLS_STR:
	case FLS_TYPE_BYTE_ARRAY: t.phys_bits = 128; break; // decoded element = 16-byte string_t
	default: t.phys_bits = 0; break;
	}
	return t;
}

// ---------------------------------------------------------------------------------------------------------
void ChunkBuilder::finish(EncodedChunk& out, fls_chunk_desc d) const {
	const std::vector<uint8_t>* segs[FLS_SEG_COUNT];
	std::vector<uint8_t> meta_bytes(meta.size() * sizeof(fls_vec_meta));
	if (!meta.empty()) {
		std::memcpy(meta_bytes.data(), meta.data(), meta_bytes.size());
	}
	segs[FLS_SEG_META] = &meta_bytes;
	segs[FLS_SEG_DATA] = &data;
	segs[FLS_SEG_AUX0] = &aux0;
	segs[FLS_SEG_AUX1] = &aux1;
	uint64_t off = 0;
	for (int s = 0; s < FLS_SEG_COUNT; ++s) {
		d.seg_off[s] = off;
		d.seg_size[s] = segs[s]->size();
		off += (segs[s]->size() + FLS_ALIGN - 1) / FLS_ALIGN * FLS_ALIGN;
	}
	// largest per-vector blob (blobs are appended in vector order, so a blob ends where the next one starts)
	uint32_t max_blob = 0;
	for (size_t v = 0; v < meta.size(); ++v) {
		const size_t begin = (size_t)meta[v].data_off16 * 16;
		const size_t end = v + 1 < meta.size() ? (size_t)meta[v + 1].data_off16 * 16 : data.size();
		if (end > begin) {
			max_blob = std::max<uint32_t>(max_blob, (uint32_t)(end - begin));
		}
	}
	d.max_blob_bytes = max_blob;
	// 128 bytes of slack so that 16-byte vector loads of the last aux elements never leave the chunk
	off += FLS_ALIGN;
	d.size = off;
	out.bytes.assign(off, 0);
	for (int s = 0; s < FLS_SEG_COUNT; ++s) {
		if (!segs[s]->empty()) {
			std::memcpy(out.bytes.data() + d.seg_off[s], segs[s]->data(), segs[s]->size());
		}
	}
	out.desc = d;
}

// ---------------------------------------------------------------------------------------------------------
template <class U>
static inline U ord_flip(bool is_signed) {
	return is_signed ? (U)((U)1 << (sizeof(U) * 8 - 1)) : (U)0;
}

template <class U>
struct VecStats {
	U omin, omax; // in order domain (sign bit flipped for signed types)
	uint32_t n_runs;
};

template <class U>
static void vec_stats(const U* v, U flip, VecStats<U>& s) {
	U mn = (U)(v[0] ^ flip), mx = mn;
	uint32_t runs = 1;
	for (int i = 1; i < kVec; ++i) {
		const U o = (U)(v[i] ^ flip);
		mn = std::min(mn, o);
		mx = std::max(mx, o);
		runs += v[i] != v[i - 1];
	}
	s.omin = mn;
	s.omax = mx;
	s.n_runs = runs;
}

// ---- FFOR (+ patches: SLPATCH, FREQUENCY) ------------------------------------------------------------------
template <class U>
static void append_packed(ChunkBuilder& b, const U* reduced, unsigned W) {
	constexpr int T = sizeof(U) * 8, LANES = kVec / T;
	const size_t old = b.data.size();
	b.data.resize(old + (size_t)W * LANES * sizeof(U));
	if (W) {
		pack_vector<U>(reduced, W, reinterpret_cast<U*>(b.data.data() + old));
	}
}

template <class U>
void encode_ffor_family(const IntChunk<U>& c, int enc, EncodedChunk& out) {
	constexpr int T = sizeof(U) * 8;
	const U flip = ord_flip<U>(c.is_signed);
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(enc);
	uint32_t n_exc_total = 0;
	unsigned maxw = 0;
	U reduced[kVec];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const U* in = c.vals + (size_t)v * kVec;
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.data_off16 = (uint32_t)(b.data.size() / 16);
		VecExceptions<U> exc;
		U base;
		unsigned W;
		if (enc == FLS_ENC_FREQUENCY) {
			base = c.frequent;
			W = 0;
			for (int i = 0; i < kVec; ++i) {
				if (in[i] != base) {
					exc.add(in[i], i);
				}
			}
		} else {
			VecStats<U> s;
			vec_stats(in, flip, s);
			base = (U)(s.omin ^ flip);
			W = bits_needed<U>((U)(s.omax - s.omin));
			if (enc == FLS_ENC_SLPATCH) {
				uint32_t hist[T + 1] = {0};
				for (int i = 0; i < kVec; ++i) {
					++hist[bits_needed<U>((U)(in[i] - base))];
				}
				uint64_t best = ~0ull;
				unsigned bestw = W;
				uint32_t above = kVec;
				for (unsigned w = 0; w <= (unsigned)T; ++w) {
					above -= hist[w];
					const uint64_t cost = 128ull * w + (uint64_t)above * (sizeof(U) + 2);
					if (cost < best) {
						best = cost;
						bestw = w;
					}
				}
				W = bestw;
			}
			const U mask = W == (unsigned)T ? (U)~(U)0 : (U)(((U)1 << W) - 1);
			for (int i = 0; i < kVec; ++i) {
				const U r = (U)(in[i] - base);
				if (r & (U)~mask) {
					exc.add(in[i], i);
					reduced[i] = 0;
				} else {
					reduced[i] = r;
				}
			}
		}
		m.base = (uint64_t)base;
		m.width = (uint8_t)W;
		m.aux_cnt = (uint16_t)exc.vals.size();
		append_packed<U>(b, reduced, W);
		exc.append_to(b);
		n_exc_total += m.aux_cnt;
		maxw = std::max(maxw, W);
		b.meta.push_back(m);
	}
	d.aux_count = n_exc_total;
	d.max_width = (uint8_t)maxw;
	b.finish(out, d);
}

// ---- DELTA -------------------------------------------------------------------------------------------------
// Transposes `in` (logical order), emits [lane bases][FFOR-packed within-lane deltas] and returns (W, dmin).
template <class U>
void delta_encode_vector(const U* in, ChunkBuilder& b, fls_vec_meta& m) {
	constexpr int T = sizeof(U) * 8, LANES = kVec / T;
	using S = typename std::make_signed<U>::type;
	U t[kVec]; // transposed, indexed by position
	for (int pos = 0; pos < kVec; ++pos) {
		t[pos] = in[fl_transpose(pos)];
	}
	U d[kVec]; // deltas indexed by position
	S dmin = 0, dmax = 0;
	bool first = true;
	for (int lane = 0; lane < LANES; ++lane) {
		for (int row = 1; row < T; ++row) {
			const U dv = (U)(t[fl_index(row, lane)] - t[fl_index(row - 1, lane)]);
			d[fl_index(row, lane)] = dv;
			const S sv = (S)dv;
			if (first) {
				dmin = dmax = sv;
				first = false;
			} else {
				dmin = std::min(dmin, sv);
				dmax = std::max(dmax, sv);
			}
		}
	}
	// row 0 stores delta == dmin so that its reduced value is 0: base[lane] = t(0,lane) - dmin
	U bases[LANES];
	for (int lane = 0; lane < LANES; ++lane) {
		bases[lane] = (U)(t[fl_index(0, lane)] - (U)dmin);
		d[fl_index(0, lane)] = (U)dmin;
	}
	const unsigned W = bits_needed<U>((U)((U)dmax - (U)dmin));
	for (int i = 0; i < kVec; ++i) {
		d[i] = (U)(d[i] - (U)dmin);
	}
	m.data_off16 = (uint32_t)(b.data.size() / 16);
	m.base = (uint64_t)(U)dmin;
	m.width = (uint8_t)W;
	// Arithmetic sequence (fls_format.h FLS_VEC_FORM_SEQUENCE): every delta equals dmin and the lane chains continue each
	// other, i.e. value(i) = value(0) + i * dmin for all 1024 positions (arithmetic modulo 2^T).  The blob is written as
	// for any DELTA vector; the mark lets a reader hand the vector on as (start, increment) without decoding it.
	if (W == 0) {
		bool seq = true;
		for (int i = 1; i < kVec && seq; ++i) {
			seq = in[i] == (U)(in[0] + (U)((U)i * (U)dmin));
		}
		if (seq) {
			m.e = FLS_VEC_FORM_SEQUENCE;
			m.reserved = (uint64_t)in[0];
		}
	}
	const size_t old = b.data.size();
	b.data.resize(old + sizeof(bases));
	std::memcpy(b.data.data() + old, bases, sizeof(bases));
	append_packed<U>(b, d, W);
}

template <class U>
void encode_delta(const IntChunk<U>& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_DELTA);
	unsigned maxw = 0;
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		delta_encode_vector<U>(c.vals + (size_t)v * kVec, b, m);
		maxw = std::max<unsigned>(maxw, m.width);
		b.meta.push_back(m);
	}
	d.max_width = (uint8_t)maxw;
	b.finish(out, d);
}

template void delta_encode_vector<uint16_t>(const uint16_t*, ChunkBuilder&, fls_vec_meta&);

#define FLS_INST(U)                                                                                                \
	template void encode_ffor_family<U>(const IntChunk<U>&, int, EncodedChunk&);                                   \
	template void encode_delta<U>(const IntChunk<U>&, EncodedChunk&);
FLS_INST(uint8_t)
FLS_INST(uint16_t)
FLS_INST(uint32_t)
FLS_INST(uint64_t)
#undef FLS_INST

} // namespace flshost
