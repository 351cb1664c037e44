// fls_encoder2.cpp — DICT / RLE / CROSS_RLE / UNCOMPRESSED / CONSTANT and the encoding selection.
#include "fls_encoder_internal.hpp"

#include <algorithm>
#include <cstring>

namespace flshost {

template <class U>
static inline U flip_of(bool is_signed) {
	return is_signed ? (U)((U)1 << (sizeof(U) * 8 - 1)) : (U)0;
}

// ---- distinct values (open addressing, early abort) -----------------------------------------------------------
template <class U>
struct DistinctSet {
	std::vector<U> keys;
	std::vector<uint32_t> counts;
	std::vector<uint8_t> used;
	uint32_t mask = 0, n = 0;
	explicit DistinctSet(uint32_t cap_pow2) : keys(cap_pow2), counts(cap_pow2, 0), used(cap_pow2, 0), mask(cap_pow2 - 1) {
	}
	static inline uint32_t hash(U k) {
		uint64_t x = (uint64_t)k * 0x9E3779B97F4A7C15ull;
		return (uint32_t)(x >> 32);
	}
	inline uint32_t slot_of(U k) const {
		uint32_t s = hash(k) & mask;
		while (used[s] && keys[s] != k) {
			s = (s + 1) & mask;
		}
		return s;
	}
	inline void add(U k) {
		const uint32_t s = slot_of(k);
		if (!used[s]) {
			used[s] = 1;
			keys[s] = k;
			++n;
		}
		++counts[s];
	}
};

template <class U>
bool encode_dict(const IntChunk<U>& c, EncodedChunk& out) {
	const size_t total = (size_t)c.n_vec * kVec;
	uint32_t cap = 1024;
	while (cap < 2 * total && cap < (1u << 30)) {
		cap <<= 1;
	}
	DistinctSet<U> set(cap);
	for (size_t i = 0; i < total; ++i) {
		set.add(c.vals[i]);
	}
	const U flip = flip_of<U>(c.is_signed);
	std::vector<U> dict;
	dict.reserve(set.n);
	for (uint32_t s = 0; s <= set.mask; ++s) {
		if (set.used[s]) {
			dict.push_back((U)(set.keys[s] ^ flip));
		}
	}
	std::sort(dict.begin(), dict.end());
	for (auto& k : dict) {
		k = (U)(k ^ flip);
	}
	// code of each key: reuse the hash table's count slots
	for (uint32_t i = 0; i < dict.size(); ++i) {
		set.counts[set.slot_of(dict[i])] = i;
	}
	const unsigned idx_bits = dict.size() <= 256 ? 8 : dict.size() <= 65536 ? 16 : 32;

	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_DICT);
	d.idx_bits = (uint8_t)idx_bits;
	d.aux_count = (uint32_t)dict.size();
	b.aux0.resize(dict.size() * sizeof(U));
	std::memcpy(b.aux0.data(), dict.data(), b.aux0.size());
	unsigned maxw = 0;
	std::vector<uint32_t> codes(kVec);
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const U* in = c.vals + (size_t)v * kVec;
		uint32_t cmin = ~0u, cmax = 0;
		for (int i = 0; i < kVec; ++i) {
			codes[i] = set.counts[set.slot_of(in[i])];
			cmin = std::min(cmin, codes[i]);
			cmax = std::max(cmax, codes[i]);
		}
		const unsigned W = bits_needed<uint32_t>(cmax - cmin);
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.data_off16 = (uint32_t)(b.data.size() / 16);
		m.base = cmin;
		m.width = (uint8_t)W;
		const size_t old = b.data.size();
		b.data.resize(old + (size_t)W * 128);
		if (W) {
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
		maxw = std::max(maxw, W);
		b.meta.push_back(m);
	}
	d.max_width = (uint8_t)maxw;
	b.finish(out, d);
	return true;
}

// ---- RLE (FastLanes style: run values + DELTA-coded u16 run index per vector) ---------------------------------
template <class U>
bool encode_rle(const IntChunk<U>& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_RLE);
	d.idx_bits = 16;
	uint32_t total_runs = 0;
	unsigned maxw = 0;
	uint16_t idxs[kVec];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const U* in = c.vals + (size_t)v * kVec;
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.aux_off = total_runs;
		uint16_t run = 0;
		auto push_val = [&](U val) {
			const size_t o = b.aux0.size();
			b.aux0.resize(o + sizeof(U));
			std::memcpy(b.aux0.data() + o, &val, sizeof(U));
		};
		push_val(in[0]);
		idxs[0] = 0;
		for (int i = 1; i < kVec; ++i) {
			if (in[i] != in[i - 1]) {
				++run;
				push_val(in[i]);
			}
			idxs[i] = run;
		}
		m.aux_cnt = (uint16_t)(run + 1u); // 1024 runs still fit: run <= 1023
		delta_encode_vector<uint16_t>(idxs, b, m);
		total_runs += (uint32_t)run + 1u;
		maxw = std::max<unsigned>(maxw, m.width);
		b.meta.push_back(m);
	}
	d.aux_count = total_runs;
	d.max_width = (uint8_t)maxw;
	b.finish(out, d);
	return true;
}

// ---- CROSS_RLE (row-group wide classic runs) -------------------------------------------------------------------
template <class U>
void encode_cross_rle(const IntChunk<U>& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_CROSS_RLE);
	const size_t total = (size_t)c.n_vec * kVec;
	std::vector<U> vals;
	std::vector<uint32_t> lens;
	size_t i = 0;
	while (i < total) {
		size_t j = i + 1;
		while (j < total && c.vals[j] == c.vals[i]) {
			++j;
		}
		vals.push_back(c.vals[i]);
		lens.push_back((uint32_t)(j - i));
		i = j;
	}
	b.aux0.resize(vals.size() * sizeof(U));
	std::memcpy(b.aux0.data(), vals.data(), b.aux0.size());
	b.aux1.resize(lens.size() * 4);
	std::memcpy(b.aux1.data(), lens.data(), b.aux1.size());
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		b.meta.push_back(m);
	}
	d.aux_count = (uint32_t)vals.size();
	b.finish(out, d);
}

template <class U>
void encode_uncompressed(const IntChunk<U>& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_UNCOMPRESSED);
	b.data.resize((size_t)c.n_vec * kVec * sizeof(U));
	std::memcpy(b.data.data(), c.vals, b.data.size());
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.data_off16 = (uint32_t)((size_t)v * kVec * sizeof(U) / 16);
		m.width = (uint8_t)(sizeof(U) * 8);
		b.meta.push_back(m);
	}
	d.max_width = (uint8_t)(sizeof(U) * 8);
	b.finish(out, d);
}

template <class U>
void encode_constant(const IntChunk<U>& c, EncodedChunk& out) {
	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_CONSTANT);
	d.const_val = (uint64_t)c.vals[0];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.base = (uint64_t)c.vals[0];
		b.meta.push_back(m);
	}
	b.finish(out, d);
}

#define FLS_INST(U)                                                                                                \
	template bool encode_dict<U>(const IntChunk<U>&, EncodedChunk&);                                               \
	template bool encode_rle<U>(const IntChunk<U>&, EncodedChunk&);                                                \
	template void encode_cross_rle<U>(const IntChunk<U>&, EncodedChunk&);                                          \
	template void encode_uncompressed<U>(const IntChunk<U>&, EncodedChunk&);                                       \
	template void encode_constant<U>(const IntChunk<U>&, EncodedChunk&);
FLS_INST(uint8_t)
FLS_INST(uint16_t)
FLS_INST(uint32_t)
FLS_INST(uint64_t)
#undef FLS_INST

} // namespace flshost
