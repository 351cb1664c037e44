// fls_encoder.hpp — product-side column-chunk encoder (host C++).
// Stands in for fastlanes::RowGroupWriter::WriteColumn + the library's encoding selection
// (reference src/writer/fls_writer.cpp:14-39; the policy itself lives in the absent library).
#pragma once
#include <cstdint>
#include <vector>
#include "fls_format.h"

namespace flshost {

struct EncodedChunk {
	fls_chunk_desc desc {};       // offset filled by the writer
	std::vector<uint8_t> bytes;   // chunk image (segments 128-aligned inside)
	int equal_to = -1;            // >= 0: EXP_EQUAL, the values are those of this earlier column of the same row group (no bytes)
};

struct ColumnType {
	uint8_t data_type = 0; // fls_data_type
	uint8_t phys_bits = 0;
	bool is_signed = false;
	bool is_float = false;
};

// values: n_tuples elements of the physical type (phys_bits wide). force_encoding < 0: smallest wins.
// Returns false and sets err if the forced encoding cannot represent the data.
bool encode_chunk(const ColumnType& type, const void* values, uint32_t n_tuples, int force_encoding,
                  EncodedChunk& out, const char** err);

// String columns (VARCHAR / BLOB): string i of the chunk = bytes[offsets[i], offsets[i+1]).  force_encoding: one of the
// string encodings (FLS_ENC_DICT_STR ...) or < 0.
bool encode_string_chunk(const uint64_t* offsets, const uint8_t* bytes, uint32_t n_tuples, int force_encoding, EncodedChunk& out,
                         const char** err);

ColumnType column_type_of(int data_type, int decimal_precision);

} // namespace flshost
