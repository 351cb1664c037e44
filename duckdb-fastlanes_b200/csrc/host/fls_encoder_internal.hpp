// fls_encoder_internal.hpp — shared pieces of the encoder translation units.
#pragma once
#include <cstdint>
#include <cstring>
#include <type_traits>
#include <vector>
#include "fls_encoder.hpp"
#include "fls_layout.hpp"

namespace flshost {

struct ChunkBuilder {
	std::vector<fls_vec_meta> meta;
	std::vector<uint8_t> data, aux0, aux1;
	void finish(EncodedChunk& out, fls_chunk_desc d) const;
};

// Exceptions of ONE vector; appended behind the vector's packed data (see fls_format.h "exception tail").
template <class U>
struct VecExceptions {
	std::vector<U> vals;
	std::vector<uint16_t> pos;
	void add(U v, int p) {
		vals.push_back(v);
		pos.push_back((uint16_t)p);
	}
	void append_to(ChunkBuilder& b) const {
		if (vals.empty()) {
			return;
		}
		const size_t n = vals.size(), o = b.data.size();
		const size_t vb = FLS_PAD16(n * sizeof(U)), pb = FLS_PAD16(n * 2);
		b.data.resize(o + vb + pb, 0);
		std::memcpy(b.data.data() + o, vals.data(), n * sizeof(U));
		std::memcpy(b.data.data() + o + vb, pos.data(), n * 2);
	}
};

// A column chunk viewed as unsigned T-bit words, padded to whole vectors with its last value.
template <class U>
struct IntChunk {
	const U* vals = nullptr; // n_vec * 1024
	uint32_t n_vec = 0;
	uint32_t n_tuples = 0;
	bool is_signed = false;
	bool is_float = false;
	U frequent = 0;          // FREQUENCY: the value to elide
	uint64_t min_val = 0, max_val = 0;
	bool has_stats = false;

	uint16_t valid_in(uint32_t v) const {
		const uint32_t left = n_tuples - v * kVec;
		return (uint16_t)(left < (uint32_t)kVec ? left : (uint32_t)kVec);
	}
	fls_chunk_desc base_desc(int enc) const {
		fls_chunk_desc d {};
		d.encoding = (uint8_t)enc;
		d.phys_bits = (uint8_t)(sizeof(U) * 8);
		d.n_vec = n_vec;
		d.n_tuples = n_tuples;
		d.min_val = min_val;
		d.max_val = max_val;
		d.has_stats = has_stats;
		return d;
	}
};

template <class U>
void encode_ffor_family(const IntChunk<U>& c, int enc, EncodedChunk& out); // FFOR, SLPATCH, FREQUENCY
template <class U>
void encode_delta(const IntChunk<U>& c, EncodedChunk& out);
template <class U>
void delta_encode_vector(const U* in, ChunkBuilder& b, fls_vec_meta& m);
template <class U>
bool encode_dict(const IntChunk<U>& c, EncodedChunk& out);      // false: too many distinct values
template <class U>
bool encode_rle(const IntChunk<U>& c, EncodedChunk& out);
template <class U>
void encode_cross_rle(const IntChunk<U>& c, EncodedChunk& out);
template <class U>
void encode_uncompressed(const IntChunk<U>& c, EncodedChunk& out);
template <class U>
void encode_constant(const IntChunk<U>& c, EncodedChunk& out);
template <class U>
void encode_alp(const IntChunk<U>& c, EncodedChunk& out);       // U = uint32_t (float) / uint64_t (double)
template <class U>
void encode_alp_rd(const IntChunk<U>& c, EncodedChunk& out);

} // namespace flshost
