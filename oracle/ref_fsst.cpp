// ref_fsst.cpp — thin extern "C" shim over the FSST implementation that ships INSIDE the reference tree
// (/root/reference/duckdb/third_party/fsst/{fsst.h,libfsst.cpp,libfsst.hpp}; the reference's fsst_kernel.hpp:54 calls the
// same fsst_decompress through the absent FastLanes library, which vendors the same cwida/fsst code).  Compiled by
// oracle/Makefile from the sources where they lie (never copied) into oracle/_ref/libfsstref.so.  TEST INFRASTRUCTURE:
// used by oracle/make_golden.py to produce tests/golden/fsst_golden.npz and by tests/test_golden.py (when the .so is
// present) to check (1) the oracle's decoder on streams the reference compressor produced and (2) the reference's
// decoder on streams the product encoder produced.
#include <cstring>
#include <vector>
#include "fsst.h"

extern "C" {

// Train on n strings (string i = bytes[off[i], off[i+1])), compress them all.  Outputs: the decoder table as
// symbol[256] (u64, little endian) + len[256], the code streams back to back and their end offsets.
// Returns the number of code bytes, or -1 if out_cap was too small.
long long ref_fsst_compress(size_t n, const unsigned long long* off, const unsigned char* bytes, unsigned long long* symbol,
                            unsigned char* len, unsigned char* out, size_t out_cap, unsigned long long* comp_end) {
	std::vector<size_t> len_in(n), len_out(n);
	std::vector<unsigned char*> str_in(n), str_out(n);
	for (size_t i = 0; i < n; ++i) {
		len_in[i] = (size_t)(off[i + 1] - off[i]);
		str_in[i] = const_cast<unsigned char*>(bytes + off[i]);
	}
	duckdb_fsst_encoder_t* enc = duckdb_fsst_create(n, len_in.data(), str_in.data(), 0);
	const duckdb_fsst_decoder_t dec = duckdb_fsst_decoder(enc);
	std::memset(symbol, 0, 256 * sizeof(unsigned long long));
	std::memset(len, 0, 256);
	for (int c = 0; c < 255; ++c) {
		symbol[c] = dec.symbol[c];
		len[c] = dec.len[c];
	}
	const size_t done = duckdb_fsst_compress(enc, n, len_in.data(), str_in.data(), out_cap, out, len_out.data(), str_out.data());
	duckdb_fsst_destroy(enc);
	if (done != n) {
		return -1;
	}
	unsigned long long end = 0;
	for (size_t i = 0; i < n; ++i) {
		if (str_out[i] != out + end) {
			return -2; // strings are expected back to back
		}
		end += len_out[i];
		comp_end[i] = end;
	}
	return (long long)end;
}

// fsst.h:162-230 with a decoder built from (symbol, len)
size_t ref_fsst_decompress(const unsigned long long* symbol, const unsigned char* len, size_t len_in, const unsigned char* in,
                           size_t out_cap, unsigned char* out) {
	duckdb_fsst_decoder_t dec;
	std::memset(&dec, 0, sizeof(dec));
	for (int c = 0; c < 255; ++c) {
		dec.symbol[c] = symbol[c];
		dec.len[c] = len[c];
	}
	return duckdb_fsst_decompress(&dec, len_in, in, out_cap, out);
}

}
