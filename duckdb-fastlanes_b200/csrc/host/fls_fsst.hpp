// fls_fsst.hpp — product-side FSST symbol-table construction and compression (host C++).
//
// The reference gets FSST from the absent FastLanes library (which vendors cwida/fsst); the code format it decodes is the
// one of duckdb/third_party/fsst/fsst.h:162-230, which IS in the reference tree: one byte per code, code c < 255 stands
// for symbol[c] (len[c] = 1..8 bytes), code 255 escapes the next byte as a literal.  Any table + code stream obeying
// that format is decodable by the reference's fsst_decompress; how the table is chosen is encoder policy.
// This file is a from-scratch trainer (iterative: encode a sample with the current table, count symbols and adjacent
// symbol pairs, keep the 255 candidates with the largest gain = frequency x length) and a greedy longest-match encoder.
#pragma once
#include <cstdint>
#include <vector>
#include "fls_format.h"

namespace flshost {

// longest-match lookup shared by both encoders: 8-byte symbols in a hash of the whole word, symbols of 2..7 bytes grouped by
// their first two bytes, longest first
struct FsstPrefixIndex {
	std::vector<uint16_t> full;   // open addressing, code + 1 (0 = empty)
	uint32_t full_shift = 60;
	std::vector<uint32_t> bucket; // 65536 entries: (first entry in `order`) << 16 | entries
	std::vector<uint16_t> order;  // codes
	uint8_t max_len = 0;

	void build(const uint64_t* symbol, const uint8_t* len, uint32_t first_code, uint32_t n_codes);
	// code of the longest symbol (>= 2 bytes) that matches at s, or -1
	int find(const uint64_t* symbol, const uint8_t* len, const uint8_t* s, uint64_t remaining, uint32_t* len_out) const;
};

struct FsstEncoder {
	fls_fsst_table table {};
	uint32_t n_symbols = 0;
	FsstPrefixIndex index;
	// single-byte symbols: direct table (code + 1)
	uint16_t byte_code[256] = {};

	// strings: n strings, string i = bytes[offsets[i] .. offsets[i+1])
	void train(const uint64_t* offsets, const uint8_t* bytes, uint64_t n);
	// appends the code stream of one string to out
	void compress(const uint8_t* s, uint64_t len, std::vector<uint8_t>& out) const;

  private:
	void rebuild_lookup();
	int find(const uint8_t* s, uint64_t remaining, uint32_t* len_out) const;
};

// FSST12 (fls_format.h fls_fsst12_table): 12-bit codes, codes < 256 are the bytes themselves, up to 3840 trained symbols of
// 2..8 bytes.  Same trainer idea (count symbols and adjacent pairs on a sample, keep the best), greedy longest match.
struct Fsst12Encoder {
	fls_fsst12_table table {};
	uint32_t n_symbols = 0; // trained symbols: codes 256 .. 256 + n_symbols - 1
	FsstPrefixIndex index;

	void train(const uint64_t* offsets, const uint8_t* bytes, uint64_t n);
	// appends the packed 12-bit code stream of one string to out
	void compress(const uint8_t* s, uint64_t len, std::vector<uint8_t>& out) const;

  private:
	void rebuild_lookup();
	int find(const uint8_t* s, uint64_t remaining, uint32_t* len_out) const; // code >= 256 or -1
};
// decoded length of a packed 12-bit code stream; ~0ull if malformed (1 trailing byte)
uint64_t fsst12_decoded_length(const fls_fsst12_table& table, const uint8_t* codes, uint64_t n);

// decoded length of a code stream under `table` (len of each code, 1 per escaped literal); ~0ull if malformed
uint64_t fsst_decoded_length(const fls_fsst_table& table, const uint8_t* codes, uint64_t n);

} // namespace flshost
