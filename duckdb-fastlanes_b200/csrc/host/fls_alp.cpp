// fls_alp.cpp — ALP and ALP-RD encoders (host C++; compile with -ffp-contract=off).
// Arithmetic follows the ALP paper (SIGMOD 2024) as stated by DuckDB's vendored copy, which IS present in the
// reference tree: duckdb/src/include/duckdb/storage/compression/alp/algorithm/alp.hpp:100-134 (encode/decode of one
// value), alp_constants.hpp:43-131 (tables), alprd/algorithm/alprd.hpp:60-236 (ALP-RD).  The sampling strategy is
// the paper's two-level one (row-group top-k, then per vector), simplified.
#include "fls_encoder_internal.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <unordered_map>

namespace flshost {

static const int64_t kFact[19] = {1ll,
                                  10ll,
                                  100ll,
                                  1000ll,
                                  10000ll,
                                  100000ll,
                                  1000000ll,
                                  10000000ll,
                                  100000000ll,
                                  1000000000ll,
                                  10000000000ll,
                                  100000000000ll,
                                  1000000000000ll,
                                  10000000000000ll,
                                  100000000000000ll,
                                  1000000000000000ll,
                                  10000000000000000ll,
                                  100000000000000000ll,
                                  1000000000000000000ll};
static const double kUpper = 9223372036854774784.0, kLower = -9223372036854774784.0;

template <class U>
struct AlpT;
template <>
struct AlpT<uint64_t> {
	using F = double;
	using I = int64_t;
	static constexpr int kMaxExp = 18;
	static constexpr double kMagic = 6755399441055744.0;
	static const double kExp[19], kFrac[19];
};
const double AlpT<uint64_t>::kExp[19] = {1.0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,
                                          1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18};
const double AlpT<uint64_t>::kFrac[19] = {1.0,   1e-1,  1e-2,  1e-3,  1e-4,  1e-5,  1e-6,  1e-7,  1e-8,  1e-9,
                                           1e-10, 1e-11, 1e-12, 1e-13, 1e-14, 1e-15, 1e-16, 1e-17, 1e-18};
template <>
struct AlpT<uint32_t> {
	using F = float;
	using I = int32_t;
	static constexpr int kMaxExp = 10;
	static constexpr float kMagic = 12582912.0f;
	static const float kExp[11], kFrac[11];
};
const float AlpT<uint32_t>::kExp[11] = {1.0F,       10.0F,       100.0F,       1000.0F,       10000.0F,      100000.0F,
                                         1000000.0F, 10000000.0F, 100000000.0F, 1000000000.0F, 10000000000.0F};
const float AlpT<uint32_t>::kFrac[11] = {1.0F,      0.1F,       0.01F,       0.001F,       0.0001F,      0.00001F,
                                          0.000001F, 0.0000001F, 0.00000001F, 0.000000001F, 0.0000000001F};

template <class U>
struct Alp {
	using A = AlpT<U>;
	using F = typename A::F;
	using I = typename A::I;

	static inline F as_f(U u) {
		F f;
		std::memcpy(&f, &u, sizeof(U));
		return f;
	}
	static inline U as_u(F f) {
		U u;
		std::memcpy(&u, &f, sizeof(U));
		return u;
	}
	static inline bool impossible(F n) {
		const double dn = (double)n;
		return !std::isfinite(n) || dn > kUpper || dn < kLower || (n == (F)0 && std::signbit(n));
	}
	// alp.hpp:108-123
	static inline int64_t encode(F value, int e, int f) {
		F tmp = value * A::kExp[e] * A::kFrac[f];
		if (impossible(tmp)) {
			return (int64_t)kUpper;
		}
		tmp = tmp + A::kMagic - A::kMagic;
		return (int64_t)tmp;
	}
	// alp.hpp:129-134
	static inline F decode(int64_t d, int e, int f) {
		return (F)d * (F)kFact[f] * A::kFrac[e];
	}
	static inline bool fits(int64_t d) {
		return d >= (int64_t)std::numeric_limits<I>::min() && d <= (int64_t)std::numeric_limits<I>::max();
	}

	struct Combo {
		int e, f;
	};

	static uint64_t estimate(const U* sample, int n, int e, int f) {
		int64_t mn = 0, mx = 0;
		int n_exc = 0, n_ok = 0;
		for (int i = 0; i < n; ++i) {
			const F x = as_f(sample[i]);
			const int64_t d = encode(x, e, f);
			if (!fits(d) || as_u(decode(d, e, f)) != sample[i]) {
				++n_exc;
				continue;
			}
			if (!n_ok++) {
				mn = mx = d;
			} else {
				mn = std::min(mn, d);
				mx = std::max(mx, d);
			}
		}
		const uint64_t range = (uint64_t)mx - (uint64_t)mn;
		const unsigned bits = range ? 64u - (unsigned)__builtin_clzll(range) : 0u;
		return (uint64_t)n * bits + (uint64_t)n_exc * (sizeof(U) * 8 + 16);
	}

	static Combo best_of_all(const U* sample, int n, uint64_t* best_size) {
		Combo best {0, 0};
		uint64_t bs = ~0ull;
		for (int e = A::kMaxExp; e >= 0; --e) {
			for (int f = e; f >= 0; --f) {
				const uint64_t s = estimate(sample, n, e, f);
				if (s < bs) {
					bs = s;
					best = {e, f};
				}
			}
		}
		if (best_size) {
			*best_size = bs;
		}
		return best;
	}
};

static inline void gather_sample(const void* vec, size_t elem, int n, void* out) {
	const int stride = kVec / n;
	for (int i = 0; i < n; ++i) {
		std::memcpy((uint8_t*)out + (size_t)i * elem, (const uint8_t*)vec + (size_t)i * stride * elem, elem);
	}
}

template <class U>
void encode_alp(const IntChunk<U>& c, EncodedChunk& out) {
	using AL = Alp<U>;
	using Combo = typename AL::Combo;
	using I = typename AL::I;
	constexpr int kSample = 32, kTopK = 5;

	// level 1: top-k (e,f) combinations over up to 8 sampled vectors of the row group
	std::vector<std::pair<int, Combo>> votes; // (count, combo)
	const uint32_t n_sv = std::min<uint32_t>(8, c.n_vec);
	U sample[kSample];
	for (uint32_t s = 0; s < n_sv; ++s) {
		const uint32_t v = (uint32_t)((uint64_t)s * c.n_vec / n_sv);
		gather_sample(c.vals + (size_t)v * kVec, sizeof(U), kSample, sample);
		const Combo cb = AL::best_of_all(sample, kSample, nullptr);
		bool found = false;
		for (auto& vt : votes) {
			if (vt.second.e == cb.e && vt.second.f == cb.f) {
				++vt.first;
				found = true;
			}
		}
		if (!found) {
			votes.push_back({1, cb});
		}
	}
	std::stable_sort(votes.begin(), votes.end(), [](const auto& a, const auto& b) { return a.first > b.first; });
	if (votes.size() > kTopK) {
		votes.resize(kTopK);
	}

	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_ALP);
	uint32_t n_exc_total = 0;
	unsigned maxw = 0;
	std::vector<int64_t> digits(kVec);
	std::vector<uint8_t> is_exc(kVec);
	U reduced[kVec];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const U* in = c.vals + (size_t)v * kVec;
		Combo cb = votes[0].second;
		if (votes.size() > 1) { // level 2: pick among the top-k on a sample of this vector
			gather_sample(in, sizeof(U), kSample, sample);
			uint64_t bs = ~0ull;
			for (const auto& vt : votes) {
				const uint64_t s = AL::estimate(sample, kSample, vt.second.e, vt.second.f);
				if (s < bs) {
					bs = s;
					cb = vt.second;
				}
			}
		}
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		m.e = (uint8_t)cb.e;
		m.f = (uint8_t)cb.f;
		VecExceptions<U> exc;
		int64_t mn = 0, mx = 0, filler = 0;
		int n_ok = 0;
		for (int i = 0; i < kVec; ++i) {
			const int64_t dg = AL::encode(AL::as_f(in[i]), cb.e, cb.f);
			const bool ok = AL::fits(dg) && AL::as_u(AL::decode(dg, cb.e, cb.f)) == in[i];
			is_exc[i] = !ok;
			digits[i] = dg;
			if (ok) {
				if (!n_ok++) {
					mn = mx = filler = dg;
				} else {
					mn = std::min(mn, dg);
					mx = std::max(mx, dg);
				}
			}
		}
		for (int i = 0; i < kVec; ++i) {
			if (is_exc[i]) {
				exc.add(in[i], i);
				digits[i] = filler;
			}
		}
		const U base = (U)(I)mn;
		const U range = (U)((U)(I)mx - base);
		const unsigned W = range ? 64u - (unsigned)__builtin_clzll((unsigned long long)range) : 0u;
		for (int i = 0; i < kVec; ++i) {
			reduced[i] = (U)((U)(I)digits[i] - base);
		}
		m.base = (uint64_t)base;
		m.width = (uint8_t)W;
		m.data_off16 = (uint32_t)(b.data.size() / 16);
		const size_t old = b.data.size();
		b.data.resize(old + (size_t)W * 128);
		if (W) {
			pack_vector<U>(reduced, W, reinterpret_cast<U*>(b.data.data() + old));
		}
		m.aux_cnt = (uint16_t)exc.vals.size();
		exc.append_to(b);
		n_exc_total += m.aux_cnt;
		maxw = std::max(maxw, W);
		b.meta.push_back(m);
	}
	d.aux_count = n_exc_total;
	d.max_width = (uint8_t)maxw;
	b.finish(out, d);
}

// ---- ALP-RD -----------------------------------------------------------------------------------------------------
template <class U>
void encode_alp_rd(const IntChunk<U>& c, EncodedChunk& out) {
	constexpr int T = sizeof(U) * 8;
	// sample the row group (alprd.hpp FindBestDictionary): cut position with the smallest estimated size
	std::vector<U> sample;
	const size_t total = (size_t)c.n_vec * kVec;
	const size_t stride = std::max<size_t>(1, total / 2048);
	for (size_t i = 0; i < total; i += stride) {
		sample.push_back(c.vals[i]);
	}
	unsigned best_right = T - 1;
	double best_est = 1e300;
	std::vector<std::pair<uint32_t, uint32_t>> best_sorted; // (count, left part)
	for (unsigned cut = 1; cut <= 16; ++cut) {
		const unsigned right = T - cut;
		std::unordered_map<uint32_t, uint32_t> hist;
		for (U x : sample) {
			++hist[(uint32_t)(x >> right)];
		}
		std::vector<std::pair<uint32_t, uint32_t>> sorted;
		for (auto& kv : hist) {
			sorted.push_back({kv.second, kv.first});
		}
		std::sort(sorted.begin(), sorted.end(), [](const auto& a, const auto& b2) {
			return a.first != b2.first ? a.first > b2.first : a.second < b2.second;
		});
		uint32_t exc = 0;
		for (size_t i = 8; i < sorted.size(); ++i) {
			exc += sorted[i].first;
		}
		const size_t nd = std::min<size_t>(8, sorted.size());
		const unsigned lw = std::max(1u, nd > 1 ? 32u - (unsigned)__builtin_clz((unsigned)(nd - 1)) : 0u);
		const double est = right + lw + (double)exc * (T + 16) / (double)sample.size();
		if (est <= best_est) {
			best_est = est;
			best_right = right;
			best_sorted = sorted;
		}
	}
	const size_t nd = std::min<size_t>(8, best_sorted.size());
	const unsigned lw = std::max(1u, nd > 1 ? 32u - (unsigned)__builtin_clz((unsigned)(nd - 1)) : 0u);

	ChunkBuilder b;
	fls_chunk_desc d = c.base_desc(FLS_ENC_ALP_RD);
	for (size_t i = 0; i < nd; ++i) {
		d.rd_dict[i] = (uint16_t)best_sorted[i].second;
	}
	const U rmask = (U)(((U)1 << best_right) - 1);
	uint32_t n_exc_total = 0;
	U right[kVec], left[kVec];
	for (uint32_t v = 0; v < c.n_vec; ++v) {
		const U* in = c.vals + (size_t)v * kVec;
		fls_vec_meta m {};
		m.n_valid = c.valid_in(v);
		VecExceptions<U> exc;
		for (int i = 0; i < kVec; ++i) {
			right[i] = in[i] & rmask;
			const uint32_t lp = (uint32_t)(in[i] >> best_right);
			int code = -1;
			for (size_t k = 0; k < nd; ++k) {
				if (d.rd_dict[k] == lp) {
					code = (int)k;
					break;
				}
			}
			if (code < 0) {
				exc.add(in[i], i);
				code = 0;
			}
			left[i] = (U)code;
		}
		m.width = (uint8_t)best_right;
		m.width2 = (uint8_t)lw;
		m.data_off16 = (uint32_t)(b.data.size() / 16);
		const size_t old = b.data.size();
		b.data.resize(old + (size_t)(best_right + lw) * 128);
		pack_vector<U>(right, best_right, reinterpret_cast<U*>(b.data.data() + old));
		pack_vector<U>(left, lw, reinterpret_cast<U*>(b.data.data() + old + (size_t)best_right * 128));
		m.aux_cnt = (uint16_t)exc.vals.size();
		exc.append_to(b);
		n_exc_total += m.aux_cnt;
		b.meta.push_back(m);
	}
	d.aux_count = n_exc_total;
	d.max_width = (uint8_t)(best_right + lw);
	b.finish(out, d);
}

template void encode_alp<uint32_t>(const IntChunk<uint32_t>&, EncodedChunk&);
template void encode_alp<uint64_t>(const IntChunk<uint64_t>&, EncodedChunk&);
template void encode_alp_rd<uint32_t>(const IntChunk<uint32_t>&, EncodedChunk&);
template void encode_alp_rd<uint64_t>(const IntChunk<uint64_t>&, EncodedChunk&);

} // namespace flshost
