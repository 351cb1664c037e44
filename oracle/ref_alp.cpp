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
// alprd.hpp:222-226: (left << right_bit_width) | right
uint64_t ref_alprd_glue_u64(uint16_t left, uint64_t right, int right_bw) {
	return (static_cast<uint64_t>(left) << right_bw) | right;
}
uint32_t ref_alprd_glue_u32(uint16_t left, uint32_t right, int right_bw) {
	return (static_cast<uint32_t>(left) << right_bw) | right;
}
int ref_alprd_max_dict_size() {
	return AlpRDConstants::MAX_DICTIONARY_SIZE;
}
int ref_alprd_max_dict_bw() {
	return AlpRDConstants::MAX_DICTIONARY_BIT_WIDTH;
}

} // extern "C"
