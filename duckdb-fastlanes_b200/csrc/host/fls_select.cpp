// fls_select.cpp — encode_chunk(): pad, collect statistics, pick the smallest encoding, encode.
// The reference leaves this policy to the absent FastLanes library (SURVEY.md 2.1 "Writer"); the rule here is
// "smallest estimated chunk wins, ties go to the cheaper decoder".
#include "fls_encoder_internal.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <unordered_map>

namespace flshost {

template <class U>
static inline unsigned nbits(U v) {
	return v ? 64u - (unsigned)__builtin_clzll((unsigned long long)v) : 0u;
}

template <class U>
struct ChunkStats {
	uint64_t ffor_bytes = 0, delta_bytes = 0, slpatch_bytes = 0, rle_bytes = 0, cross_rle_bytes = 0;
	uint64_t dict_bytes = ~0ull, freq_bytes = ~0ull;
	bool all_equal = true;
	U frequent = 0;
};

template <class U>
static void collect_stats(const IntChunk<U>& c, ChunkStats<U>& st) {
	constexpr int T = sizeof(U) * 8;
	using S = typename std::make_signed<U>::type;
	const U flip = c.is_signed ? (U)((U)1 << (T - 1)) : (U)0;
	const size_t total = (size_t)c.n_vec * kVec;
	uint64_t cross_runs = 1;
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const U* in = c.vals + (size_t)v * kVec;
		U mn = (U)(in[0] ^ flip), mx = mn;
		uint32_t runs = 1;
		S dmin = 0, dmax = 0;
		bool dfirst = true;
		for (int i = 1; i < kVec; ++i) {
			const U o = (U)(in[i] ^ flip);
			mn = std::min(mn, o);
			mx = std::max(mx, o);
			runs += in[i] != in[i - 1];
			if (i % T) { // chain-internal step (a lane chain covers T logically consecutive values)
				const S dv = (S)(U)(in[i] - in[i - 1]);
				if (dfirst) {
					dmin = dmax = dv;
					dfirst = false;
				} else {
					dmin = std::min(dmin, dv);
					dmax = std::max(dmax, dv);
				}
			}
		}
		if (v > 0 && in[0] != in[-1]) {
			++cross_runs;
		}
		cross_runs += runs - 1;
		const unsigned W = nbits<U>((U)(mx - mn));
		st.ffor_bytes += 128ull * W + sizeof(fls_vec_meta);
		st.delta_bytes += 128ull * nbits<U>((U)((U)dmax - (U)dmin)) + 128 + sizeof(fls_vec_meta);
		// RLE: run values + DELTA-coded u16 index (width <= 1 bit when runs are long)
		st.rle_bytes += (uint64_t)runs * sizeof(U) + 128 + 128 * (runs > 1 ? 1 : 0) + sizeof(fls_vec_meta);
		// SLPATCH
		const U base = (U)(mn ^ flip);
		uint32_t hist[T + 1];
		std::memset(hist, 0, sizeof(hist));
		for (int i = 0; i < kVec; ++i) {
			++hist[nbits<U>((U)(in[i] - base))];
		}
		uint64_t best = ~0ull;
		uint32_t above = kVec;
		for (unsigned w = 0; w <= W; ++w) {
			above -= hist[w];
			best = std::min<uint64_t>(best, 128ull * w + (uint64_t)above * (sizeof(U) + 2));
		}
		st.slpatch_bytes += best + sizeof(fls_vec_meta);
		if (mx != mn || in[0] != c.vals[0]) {
			st.all_equal = false;
		}
	}
	st.cross_rle_bytes = cross_runs * (sizeof(U) + 4) + (uint64_t)c.n_vec * sizeof(fls_vec_meta);

	// distinct values: sample first, count fully only when a dictionary could win
	const size_t stride = std::max<size_t>(1, total / 1024);
	std::unordered_map<uint64_t, uint32_t> sample;
	for (size_t i = 0; i < total; i += stride) {
		++sample[(uint64_t)c.vals[i]];
	}
	const size_t n_sampled = (total + stride - 1) / stride;
	if (sample.size() * 2 <= n_sampled || total <= 4096) {
		std::unordered_map<uint64_t, uint32_t> all;
		all.reserve(4096);
		const size_t limit = total / 4 + 2;
		bool aborted = false;
		for (size_t i = 0; i < total; ++i) {
			++all[(uint64_t)c.vals[i]];
			if (all.size() > limit) {
				aborted = true;
				break;
			}
		}
		if (!aborted) {
			const uint64_t nd = all.size();
			st.dict_bytes = nd * sizeof(U) + (uint64_t)c.n_vec * (128ull * nbits<uint64_t>(nd - 1) + sizeof(fls_vec_meta));
			uint32_t best = 0;
			for (auto& kv : all) {
				if (kv.second > best || (kv.second == best && (U)kv.first < st.frequent)) {
					best = kv.second;
					st.frequent = (U)kv.first;
				}
			}
			st.freq_bytes = (uint64_t)(total - best) * (sizeof(U) + 2) + (uint64_t)c.n_vec * sizeof(fls_vec_meta);
		}
	}
}

template <class U>
static void fill_minmax(IntChunk<U>& c, const ColumnType& type) {
	constexpr int T = sizeof(U) * 8;
	c.has_stats = true;
	if (type.is_float) {
		double mn = INFINITY, mx = -INFINITY;
		bool any = false;
		for (uint32_t i = 0; i < c.n_tuples; ++i) {
			double x;
			if (T == 32) {
				float f;
				std::memcpy(&f, &c.vals[i], 4);
				x = f;
			} else {
				std::memcpy(&x, &c.vals[i], 8);
			}
			if (x == x) {
				mn = std::min(mn, x);
				mx = std::max(mx, x);
				any = true;
			}
		}
		c.has_stats = any;
		std::memcpy(&c.min_val, &mn, 8); // zone maps of float columns are kept as doubles
		std::memcpy(&c.max_val, &mx, 8);
		return;
	}
	if (type.is_signed) {
		using S = typename std::make_signed<U>::type;
		S mn = (S)c.vals[0], mx = mn;
		for (uint32_t i = 1; i < c.n_tuples; ++i) {
			mn = std::min(mn, (S)c.vals[i]);
			mx = std::max(mx, (S)c.vals[i]);
		}
		c.min_val = (uint64_t)(int64_t)mn;
		c.max_val = (uint64_t)(int64_t)mx;
	} else {
		U mn = c.vals[0], mx = mn;
		for (uint32_t i = 1; i < c.n_tuples; ++i) {
			mn = std::min(mn, c.vals[i]);
			mx = std::max(mx, c.vals[i]);
		}
		c.min_val = mn;
		c.max_val = mx;
	}
}

template <class U>
static bool encode_typed(const ColumnType& type, const void* values, uint32_t n_tuples, int force, EncodedChunk& out,
                         const char** err) {
	const uint32_t n_vec = (n_tuples + kVec - 1) / kVec;
	std::vector<U> padded;
	const U* vals = static_cast<const U*>(values);
	if (n_tuples % kVec) {
		padded.assign(vals, vals + n_tuples);
		padded.resize((size_t)n_vec * kVec, vals[n_tuples - 1]);
		vals = padded.data();
	}
	IntChunk<U> c;
	c.vals = vals;
	c.n_vec = n_vec;
	c.n_tuples = n_tuples;
	c.is_signed = type.is_signed;
	c.is_float = type.is_float;
	fill_minmax(c, type);

	int enc = force;
	ChunkStats<U> st;
	if (enc < 0 || enc == FLS_ENC_FREQUENCY) {
		collect_stats(c, st);
		c.frequent = st.frequent;
		if (enc == FLS_ENC_FREQUENCY && st.freq_bytes == ~0ull) {
			// statistics pass gave up on counting: take the first value as the frequent one
			c.frequent = vals[0];
		}
	}
	if (enc < 0) {
		if (st.all_equal) {
			enc = FLS_ENC_CONSTANT;
		} else if (type.is_float) {
			// ALP family is tried for real (its size is data dependent), then compared with the generic candidates
			EncodedChunk alp, rd;
			if constexpr (sizeof(U) >= 4) {
				encode_alp<U>(c, alp);
				uint64_t best = alp.bytes.size();
				const uint64_t raw = (uint64_t)n_vec * kVec * sizeof(U);
				if (best * 10 > raw * 6) { // poor ALP ratio: the column holds "real" doubles
					encode_alp_rd<U>(c, rd);
					if (rd.bytes.size() < best) {
						best = rd.bytes.size();
						alp = std::move(rd);
					}
				}
				const uint64_t generic = std::min({st.dict_bytes, st.rle_bytes, st.cross_rle_bytes});
				if (best <= generic && best < raw) {
					out = std::move(alp);
					return true;
				}
				if (generic >= raw) {
					enc = FLS_ENC_UNCOMPRESSED;
				} else {
					enc = generic == st.dict_bytes ? FLS_ENC_DICT : generic == st.rle_bytes ? FLS_ENC_RLE : FLS_ENC_CROSS_RLE;
				}
			}
		} else {
			struct Cand {
				int enc;
				uint64_t bytes;
			};
			// listed from cheapest decoder to most expensive; strict '<' keeps the earlier one on ties
			const Cand cands[] = {{FLS_ENC_FFOR, st.ffor_bytes},           {FLS_ENC_DELTA, st.delta_bytes},
			                      {FLS_ENC_DICT, st.dict_bytes},           {FLS_ENC_SLPATCH, st.slpatch_bytes},
			                      {FLS_ENC_FREQUENCY, st.freq_bytes},      {FLS_ENC_RLE, st.rle_bytes},
			                      {FLS_ENC_CROSS_RLE, st.cross_rle_bytes}};
			uint64_t best = ~0ull;
			for (const auto& cd : cands) {
				// a more complex scheme must save at least 1/16 to displace a simpler one
				if (cd.bytes < best - best / 16) {
					best = cd.bytes;
					enc = cd.enc;
				}
			}
		}
	}
	switch (enc) {
	case FLS_ENC_UNCOMPRESSED: encode_uncompressed<U>(c, out); return true;
	case FLS_ENC_CONSTANT:
		for (size_t i = 1; i < (size_t)n_vec * kVec; ++i) {
			if (vals[i] != vals[0]) {
				*err = "CONSTANT forced on a non-constant chunk";
				return false;
			}
		}
		encode_constant<U>(c, out);
		return true;
	case FLS_ENC_FFOR:
	case FLS_ENC_SLPATCH:
	case FLS_ENC_FREQUENCY: encode_ffor_family<U>(c, enc, out); return true;
	case FLS_ENC_DELTA:
		// The padding of a partial last vector is never exposed (n_valid).  For DELTA, continue the last step of the data
		// instead of repeating the last value when the data ends on a constant step: a sorted / arithmetic column then
		// keeps its delta width in the tail vector (a repeated value puts a 0 delta among constant non-zero deltas, and one
		// wide vector sizes the decoder's shared-memory slots for the whole launch).
		if (!padded.empty() && n_tuples % kVec >= 3) {
			const U step = (U)(padded[n_tuples - 1] - padded[n_tuples - 2]);
			if (step != 0 && step == (U)(padded[n_tuples - 2] - padded[n_tuples - 3])) {
				for (size_t i = n_tuples; i < (size_t)n_vec * kVec; ++i) {
					padded[i] = (U)(padded[i - 1] + step);
				}
			}
		}
		encode_delta<U>(c, out);
		return true;
	case FLS_ENC_DICT: return encode_dict<U>(c, out);
	case FLS_ENC_RLE: return encode_rle<U>(c, out);
	case FLS_ENC_CROSS_RLE: encode_cross_rle<U>(c, out); return true;
	case FLS_ENC_ALP:
	case FLS_ENC_ALP_RD:
		if constexpr (sizeof(U) >= 4) {
			if (type.is_float) {
				if (enc == FLS_ENC_ALP) {
					encode_alp<U>(c, out);
				} else {
					encode_alp_rd<U>(c, out);
				}
				return true;
			}
		}
		*err = "ALP encodings need a FLOAT/DOUBLE column";
		return false;
	default: *err = "unknown encoding"; return false;
	}
}

bool encode_chunk(const ColumnType& type, const void* values, uint32_t n_tuples, int force_encoding, EncodedChunk& out,
                  const char** err) {
	static const char* none = "";
	if (!err) {
		err = &none;
	}
	if (n_tuples == 0) {
		*err = "empty column chunk";
		return false;
	}
	switch (type.phys_bits) {
	case 8: return encode_typed<uint8_t>(type, values, n_tuples, force_encoding, out, err);
	case 16: return encode_typed<uint16_t>(type, values, n_tuples, force_encoding, out, err);
	case 32: return encode_typed<uint32_t>(type, values, n_tuples, force_encoding, out, err);
	case 64: return encode_typed<uint64_t>(type, values, n_tuples, force_encoding, out, err);
	default: *err = "unsupported physical width"; return false;
	}
}

} // namespace flshost
