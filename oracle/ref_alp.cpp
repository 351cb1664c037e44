// ref_alp.cpp — thin extern "C" shim over the ALP / ALP-RD arithmetic that ships INSIDE the reference tree
// (DuckDB's vendored implementation, /root/reference/duckdb/src/include/duckdb/storage/compression/alp/algorithm/alp.hpp
// and alprd/algorithm/alprd.hpp).  Compiled by oracle/Makefile from the headers where they lie (never copied) into
// oracle/_ref/libalpref.so.  TEST INFRASTRUCTURE: used by oracle/make_golden.py to produce tests/golden/alp_*.npz and by
// tests/test_golden.py (when the .so is present) to check the oracle restatement value by value.
#include "duckdb/storage/compression/alp/algorithm/alp.hpp"
#include "duckdb/storage/compression/alprd/algorithm/alprd.hpp"

using namespace duckdb;

// The four Value:: helpers IsImpossibleToEncode calls live in duckdb/src/common/types/value.cpp:536-553, a translation
// unit that drags in the whole DuckDB library; they are one-liners over <cmath>, supplied here so the shim links alone.
namespace duckdb {
template <>
bool Value::IsNan(float input) {
	return std::isnan(input);
}
template <>
bool Value::IsNan(double input) {
	return std::isnan(input);
}
template <>
bool Value::IsFinite(float input) {
	return !(std::isnan(input) || std::isinf(input));
}
template <>
bool Value::IsFinite(double input) {
	return !(std::isnan(input) || std::isinf(input));
}
} // namespace duckdb

// AlpRDDecompression<T>::Decompress itself (alprd.hpp:203-236) over one vector: the left codes and right parts are packed
// with DuckDB's own BitpackingPrimitives (the layout Decompress unpacks), then the reference's decompression runs unchanged:
// bit-unpacking, dictionary lookup, glue, exception patching.  exc_left = the exceptions' left parts (alprd.hpp:229-234).
template <class T, class EXACT>
static void alprd_decompress(const uint16_t* left_codes, const EXACT* right_parts, int n, const uint16_t* dict, const uint16_t* exc_left,
                             const uint16_t* exc_pos, int n_exc, int left_bw, int right_bw, EXACT* out) {
	std::vector<uint8_t> left_packed(AlpRDConstants::ALP_VECTOR_SIZE * 8 + 1024, 0), right_packed(AlpRDConstants::ALP_VECTOR_SIZE * 8 + 1024, 0);
	BitpackingPrimitives::PackBuffer<uint16_t>(left_packed.data(), const_cast<uint16_t*>(left_codes), (idx_t)n, (bitpacking_width_t)left_bw);
	BitpackingPrimitives::PackBuffer<EXACT>(right_packed.data(), const_cast<EXACT*>(right_parts), (idx_t)n, (bitpacking_width_t)right_bw);
	alp::AlpRDDecompression<T>::Decompress(left_packed.data(), right_packed.data(), dict, out, (idx_t)n, (uint16_t)n_exc, exc_left, exc_pos,
	                                       (uint8_t)left_bw, (uint8_t)right_bw);
}

extern "C" {

// alp.hpp:129-134
double ref_alp_decode_f64(int64_t digit, int e, int f) {
	return alp::AlpCompression<double, true>::DecodeValue(digit, alp::AlpEncodingIndices((uint8_t)e, (uint8_t)f));
}
float ref_alp_decode_f32(int64_t digit, int e, int f) {
	return alp::AlpCompression<float, true>::DecodeValue(digit, alp::AlpEncodingIndices((uint8_t)e, (uint8_t)f));
}
// alp.hpp:108-123 (magic-number rounding; impossible values map to ENCODING_UPPER_LIMIT)
int64_t ref_alp_encode_f64(double v, int e, int f) {
	return alp::AlpCompression<double, true>::EncodeValue(v, alp::AlpEncodingIndices((uint8_t)e, (uint8_t)f));
}
int64_t ref_alp_encode_f32(float v, int e, int f) {
	return alp::AlpCompression<float, true>::EncodeValue(v, alp::AlpEncodingIndices((uint8_t)e, (uint8_t)f));
}
int ref_alp_impossible_f64(double v) {
	return alp::AlpCompression<double, true>::IsImpossibleToEncode(v);
}
int ref_alp_impossible_f32(float v) {
	return alp::AlpCompression<float, true>::IsImpossibleToEncode(v);
}
void ref_alprd_decompress_u64(const uint16_t* left_codes, const uint64_t* right_parts, int n, const uint16_t* dict, const uint16_t* exc_left,
                              const uint16_t* exc_pos, int n_exc, int left_bw, int right_bw, uint64_t* out) {
	alprd_decompress<double, uint64_t>(left_codes, right_parts, n, dict, exc_left, exc_pos, n_exc, left_bw, right_bw, out);
}
void ref_alprd_decompress_u32(const uint16_t* left_codes, const uint32_t* right_parts, int n, const uint16_t* dict, const uint16_t* exc_left,
                              const uint16_t* exc_pos, int n_exc, int left_bw, int right_bw, uint32_t* out) {
	alprd_decompress<float, uint32_t>(left_codes, right_parts, n, dict, exc_left, exc_pos, n_exc, left_bw, right_bw, out);
}
int ref_alprd_max_dict_size() {
	return AlpRDConstants::MAX_DICTIONARY_SIZE;
}
int ref_alprd_max_dict_bw() {
	return AlpRDConstants::MAX_DICTIONARY_BIT_WIDTH;
}

} // extern "C"
