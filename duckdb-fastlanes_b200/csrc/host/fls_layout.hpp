// fls_layout.hpp — FastLanes unified transposed layout, host side (product encoder).
//
// Geometry (FastLanes paper, VLDB 2023): a vector is 1024 values; for a T-bit
// physical type there are LANES = 1024/T interleaved lanes of T rows each.  The value stored at (row, lane)
// is the one at logical position fl_index(row, lane).  DELTA additionally permutes the vector with
// fl_transpose() so that every lane chain covers T logically consecutive values.
#pragma once
#include <cstdint>
#include <cstring>

namespace flshost {

static constexpr int kVec = 1024;
static constexpr int kFlOrder[8] = {0, 4, 2, 6, 1, 5, 3, 7};

inline int fl_index(int row, int lane) {
	return kFlOrder[row >> 3] * 16 + (row & 7) * 128 + lane;
}
// transposed position -> logical position
inline int fl_transpose(int pos) {
	const int lane = pos & 15, order = (pos >> 4) & 7, row = pos >> 7;
	return lane * 64 + kFlOrder[order] * 8 + row;
}

template <class U>
inline unsigned bits_needed(U v) {
	unsigned n = 0;
	while (v) {
		++n;
		v = (U)(v >> 1);
	}
	return n;
}

// Pack 1024 already-reduced values (indexed by position) at width W into W*LANES words of type U.
template <class U>
inline void pack_vector(const U* in, unsigned W, U* out) {
	constexpr int T = sizeof(U) * 8, LANES = kVec / T;
	std::memset(out, 0, (size_t)W * LANES * sizeof(U));
	if (W == 0) {
		return;
	}
	const U mask = (W == (unsigned)T) ? (U)~(U)0 : (U)(((U)1 << W) - 1);
	for (int lane = 0; lane < LANES; ++lane) {
		for (int row = 0; row < T; ++row) {
			const U v = in[fl_index(row, lane)] & mask;
			const unsigned bit = row * W, word = bit / T, sh = bit % T;
			out[word * LANES + lane] |= (U)(v << sh);
			if (sh + W > (unsigned)T) {
				out[(word + 1) * LANES + lane] |= (U)(v >> (T - sh));
			}
		}
	}
}

} // namespace flshost
