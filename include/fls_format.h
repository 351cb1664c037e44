/*
 * fls_format.h — on-disk / in-HBM container read by the B200 FastLanes decode engine.
 *
 * Why a container of our own: the reference extension (sebastiaan-dev/duckdb-fastlanes) delegates the
 * .fls byte layout to a third-party library that is not vendored (reference CMakeLists.txt:56-62), so the
 * true footer/segment bytes cannot be known here.  What IS pinned by the reference is the logical model:
 *   table -> row groups (64 vectors x 1024 values, reference src/include/writer/fls_writer.hpp:14)
 *         -> one column chunk per (row group, column), each with ONE encoding expression
 *         -> per-vector entry points into segments (reference src/reader/fls_reader.cpp:448,458: get_chunk /
 *            PointTo(vec_idx)).
 * This header states that model as plain C structs.  It is a FORMAT contract shared by the product encoder,
 * the CUDA decoder and the CPU oracle; it contains no algorithm.
 *
 * All integers little-endian.  All segment starts are 128-byte aligned relative to the file start so that a
 * staged copy in HBM (256-byte aligned base) satisfies the 16-byte alignment of TMA bulk copies.
 */
#ifndef FLS_FORMAT_H
#define FLS_FORMAT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLS_MAGIC "FLSB2001"
#define FLS_VERSION 2u
#define FLS_VEC_SZ 1024u            /* fastlanes::CFG::VEC_SZ */
#define FLS_DEFAULT_VEC_PER_RG 64u  /* row group = 65 536 tuples, reference fls_writer.hpp:14 */
#define FLS_ALIGN 128u

/* Mirrors the names of fastlanes::DataType that the reference maps (src/reader/translation_utils.cpp:5-48,
 * src/writer/translation_utils.cpp:6-70). */
typedef enum fls_data_type {
	FLS_TYPE_INVALID = 0,
	FLS_TYPE_INT8 = 1,
	FLS_TYPE_INT16 = 2,
	FLS_TYPE_INT32 = 3,
	FLS_TYPE_INT64 = 4,
	FLS_TYPE_UINT8 = 5,
	FLS_TYPE_UINT16 = 6,
	FLS_TYPE_UINT32 = 7,
	FLS_TYPE_UINT64 = 8,
	FLS_TYPE_FLOAT = 9,
	FLS_TYPE_DOUBLE = 10,
	FLS_TYPE_BOOLEAN = 11, /* u8 view, reference fls_view_writer.cpp:63-64 */
	FLS_TYPE_DATE = 12,    /* int32 days */
	FLS_TYPE_TIMESTAMP = 13, /* int64 microseconds */
	FLS_TYPE_DECIMAL = 14, /* physical INT16/32/64 chosen by precision; fix_me_decimal_type in the reference */
	FLS_TYPE_FLS_STR = 15, /* VARCHAR */
	FLS_TYPE_BYTE_ARRAY = 16
} fls_data_type;

/* One encoding expression per column chunk (the final operator of the reference's PhysicalExpr,
 * src/reader/fls_reader.cpp:459; operator structs listed in SURVEY.md section 2.2). */
typedef enum fls_encoding {
	FLS_ENC_UNCOMPRESSED = 0, /* dec_uncompressed_opr */
	FLS_ENC_CONSTANT = 1,     /* dec_constant_opr */
	FLS_ENC_FFOR = 2,         /* dec_unffor_opr: bit-unpack + frame of reference */
	FLS_ENC_DELTA = 3,        /* dec_unffor -> rsum -> dec_transpose_opr */
	FLS_ENC_DICT = 4,         /* dec_dict_opr<KEY,IDX>: codes FFOR-packed in IDX lanes */
	FLS_ENC_RLE = 5,          /* dec_rle_map_opr<KEY,u16>: run values + DELTA-coded run index */
	FLS_ENC_ALP = 6,          /* dec_alp_opr */
	FLS_ENC_ALP_RD = 7,       /* dec_alp_rd_opr */
	FLS_ENC_CROSS_RLE = 8,    /* dec_cross_rle_opr: row-group wide (value,length) runs */
	FLS_ENC_FREQUENCY = 9,    /* dec_frequency_opr: frequent value + exceptions */
	FLS_ENC_SLPATCH = 10,     /* dec_slpatch_opr: FFOR at reduced width + patches */
	FLS_ENC_DICT_STR = 11,    /* dec_dict_opr<fls_string_t,IDX>: dictionary of strings + packed codes */
	FLS_ENC_STR_UNCOMPRESSED = 12, /* dec_fls_str_uncompressed_opr: lengths + bytes */
	FLS_ENC_STR_CONSTANT = 13, /* dec_constant_str_opr */
	FLS_ENC_FSST = 14,        /* dec_fsst_opr: FSST-compressed strings (8-bit codes, 255 = escape) */
	FLS_ENC_FSST_DICT = 15,   /* dec_fsst_dict_opr<IDX>: dictionary whose strings are FSST-compressed + packed codes */
	FLS_ENC_FSST12 = 16,      /* dec_fsst12_opr: FSST with 12-bit codes (4096 symbols, codes < 256 = the byte itself, no escapes) */
	FLS_ENC_FSST12_DICT = 17, /* dec_fsst12_dict_opr<IDX> */
	FLS_ENC_RLE_STR = 18,       /* dec_rle_map_opr<fls_string_t,u16> (rle_map_kernel.hpp:26-61): run strings + DELTA-coded run index */
	FLS_ENC_CROSS_RLE_STR = 19, /* dec_cross_rle_opr<fls_string_t> (cross_rle_kernel.hpp:55-97,121-144): row-group wide (string,length) runs */
	FLS_ENC_FREQUENCY_STR = 20, /* dec_frequency_str_opr (frequency_kernel.hpp:23-69): frequent string + exceptions */
	FLS_ENC_COUNT = 21
} fls_encoding;

#define FLS_ENC_IS_FSST8(e) ((e) == FLS_ENC_FSST || (e) == FLS_ENC_FSST_DICT)
#define FLS_ENC_IS_FSST12(e) ((e) == FLS_ENC_FSST12 || (e) == FLS_ENC_FSST12_DICT)
#define FLS_ENC_IS_FSST_ANY(e) (FLS_ENC_IS_FSST8(e) || FLS_ENC_IS_FSST12(e))
#define FLS_ENC_IS_FSST_ROWS(e) ((e) == FLS_ENC_FSST || (e) == FLS_ENC_FSST12) /* one compressed string per row */
#define FLS_ENC_IS_STR_DICT(e) ((e) == FLS_ENC_DICT_STR || (e) == FLS_ENC_FSST_DICT || (e) == FLS_ENC_FSST12_DICT)
/* string encodings whose rows name one of aux_count VALUE strings (bytes in AUX0, u32 end[aux_count] in AUX1) by a rule of
 * their own: run index, run position, exception list */
#define FLS_ENC_IS_STR_VALUES(e) ((e) == FLS_ENC_RLE_STR || (e) == FLS_ENC_CROSS_RLE_STR || (e) == FLS_ENC_FREQUENCY_STR)

/* Segment slots inside a column chunk. */
enum {
	FLS_SEG_META = 0, /* fls_vec_meta[n_vec] */
	FLS_SEG_DATA = 1, /* per-vector packed blobs, addressed by fls_vec_meta.data_off16 */
	FLS_SEG_AUX0 = 2, /* DICT keys | RLE run values | CROSS_RLE values | string bytes ([FSST table] + bytes) */
	FLS_SEG_AUX1 = 3, /* CROSS_RLE lengths (u32) | string end offsets (u32; FSST: compressed ends, then decoded ends) */
	FLS_SEG_COUNT = 4
};

typedef struct fls_file_header { /* 64 bytes at offset 0 */
	char magic[8];
	uint32_t version;
	uint32_t n_columns;
	uint32_t n_rowgroups;
	uint32_t vec_per_rowgroup;
	uint64_t n_tuples;
	uint64_t footer_offset; /* fls_column_desc[n_columns], fls_rowgroup_desc[n_rowgroups],
	                           fls_chunk_desc[n_rowgroups*n_columns] (row-group major) */
	uint64_t footer_size;
	uint8_t reserved[16];
} fls_file_header;

typedef struct fls_column_desc { /* 80 bytes */
	char name[64];
	uint8_t data_type;  /* fls_data_type */
	uint8_t phys_bits;  /* 8/16/32/64: width T of the stored integers / IEEE words */
	uint8_t is_signed;  /* integer interpretation of the physical words */
	uint8_t is_float;   /* 1: FLOAT/DOUBLE payload */
	uint8_t decimal_precision;
	uint8_t decimal_scale;
	uint8_t reserved[10];
} fls_column_desc;

typedef struct fls_rowgroup_desc { /* 24 bytes */
	uint64_t first_tuple; /* file_row_number of tuple 0 of this row group */
	uint32_t n_tuples;
	uint32_t n_vec;
	uint64_t reserved;
} fls_rowgroup_desc;

typedef struct fls_chunk_desc { /* 144 bytes */
	uint64_t offset; /* from file start, 128-aligned */
	uint64_t size;
	uint64_t seg_off[FLS_SEG_COUNT]; /* from chunk start, 128-aligned */
	uint64_t seg_size[FLS_SEG_COUNT];
	uint64_t const_val; /* CONSTANT: the value (zero-extended physical word) */
	uint64_t min_val;   /* zone map: integers as the physical word sign- / zero-extended to 64 bits; FLOAT and DOUBLE columns as the bits
	                       of an IEEE double (FLOAT values widened; NaNs are left out, has_stats = 0 when the chunk holds nothing else) */
	uint64_t max_val;
	uint32_t n_vec;
	uint32_t n_tuples;
	uint32_t aux_count;  /* DICT: #keys; RLE/CROSS_RLE: total #runs; patched encodings: total #exceptions */
	uint8_t encoding;    /* fls_encoding */
	uint8_t phys_bits;   /* T */
	uint8_t idx_bits;    /* DICT: lane width of the codes (8/16/32); RLE: 16 */
	uint8_t max_width;   /* max fls_vec_meta.width(+width2) over the chunk: sizes the shared-memory stage */
	uint8_t has_stats;
	uint8_t is_equal;     /* 1: EXP_EQUAL (multi-column compression, reference fls_reader.cpp:43-82): the chunk's values are those of
	                         column equal_col's chunk of the same row group.  The descriptor is a copy of that chunk's (same offset
	                         and segments: the bytes exist once), so every decoder reads it like any other chunk */
	uint8_t equal_col[2]; /* little-endian u16 */
	uint16_t rd_dict[8]; /* ALP_RD: left-part dictionary (alprd_constants.hpp MAX_DICTIONARY_SIZE = 8) */
	uint32_t max_blob_bytes; /* largest per-vector blob of the chunk (packed data + exception tail): sizes the TMA slots */
} fls_chunk_desc;

/* Per-vector entry point + parameters: what PointTo(vec_idx) resolves in the reference. */
typedef struct fls_vec_meta { /* 32 bytes */
	uint64_t base;       /* FFOR/SLPATCH/ALP: frame of reference; FREQUENCY: frequent value */
	uint32_t data_off16; /* blob offset from the DATA segment start, 16-byte units */
	uint32_t aux_off;    /* RLE: first run value of this vector in AUX0, in elements */
	uint16_t aux_cnt;    /* #exceptions (stored in the blob's tail) | #runs of this vector */
	uint8_t width;       /* packed bit width W in [0,T] */
	uint8_t width2;      /* ALP_RD: bit width of the left-part codes */
	uint8_t e;           /* ALP exponent; DELTA: vector form mark (FLS_VEC_FORM_*) */
	uint8_t f;           /* ALP factor */
	uint16_t n_valid;    /* tuples of this vector that exist (last vector of a row group may be partial) */
	uint64_t reserved;   /* DELTA vectors marked FLS_VEC_FORM_SEQUENCE: value(0), zero-extended physical word */
} fls_vec_meta;

/* Vector forms a reader may hand on WITHOUT decoding, the way the reference hands a dec_constant_opr chunk to DuckDB as a
 * CONSTANT_VECTOR (constant_kernel.hpp:19-27).  They are facts the writer established from the data and recorded in the
 * entry point; the packed blob is complete either way, so every decoder may ignore them.
 *   CONSTANT: chunk encoding CONSTANT, or an FFOR vector of width 0 without exceptions (every value = base)
 *   SEQUENCE: a DELTA vector of width 0 whose fls_vec_meta.e == FLS_VEC_FORM_SEQUENCE:
 *             value(i) = reserved + i * base in arithmetic modulo 2^T, i = 0..1023 (DuckDB SEQUENCE_VECTOR, vector_type.hpp:20) */
#define FLS_VEC_FORM_FLAT 0u
#define FLS_VEC_FORM_CONSTANT 1u
#define FLS_VEC_FORM_SEQUENCE 2u

/* Per-vector blob layout in the DATA segment (sizes in bytes, every blob is a multiple of 16 and starts 16-aligned):
 *   FFOR / SLPATCH / ALP / FREQUENCY(W=0): [128*W packed][exception tail]
 *   ALP_RD:                                [128*W right parts][128*W2 left codes][exception tail]  (both in T lanes)
 *   DICT codes:                            [128*W packed]
 *   DELTA, RLE index:                      [128 lane bases][128*W packed deltas]
 *   UNCOMPRESSED:                          [1024*T/8 natural order]
 * exception tail (aux_cnt = n > 0 only):   [n values of T bits, padded to 16 B][n positions u16, padded to 16 B]
 * The exceptions travel with the vector so that ONE bulk copy stages everything a vector needs (out[pos[k]] = value[k]
 * after the vector has been decoded: alp.hpp:401-404, slpatch / frequency patches).
 */
/* String columns (data_type FLS_TYPE_FLS_STR / BYTE_ARRAY, phys_bits = 128: the decoded element is DuckDB's 16-byte
 * string_t).  "n" below is the number of strings the offsets cover: n_vec*1024 for per-row encodings (the padding rows
 * of a partial last vector are empty strings), aux_count (#keys) for the dictionary encodings.
 *   STR_UNCOMPRESSED: AUX0 = bytes of all strings back to back; AUX1 = u32 end[n].                 const_val = bytes
 *   STR_CONSTANT:     AUX0 = the string; aux_count = its length.                                  const_val = length
 *   DICT_STR:         META/DATA = packed codes as for DICT (idx_bits lanes); AUX0 = dictionary bytes; AUX1 = u32 end[n_keys]
 *   FSST:             AUX0 = fls_fsst_table | compressed bytes; AUX1 = u32 comp_end[n] | u32 dec_end[n]
 *   FSST_DICT:        META/DATA = packed codes; AUX0 = fls_fsst_table | compressed dictionary; AUX1 = comp_end[n_keys] | dec_end[n_keys]
 *   FSST12 / FSST12_DICT: the same with fls_fsst12_table and 12-bit code streams
 *   RLE_STR:          per vector the strings of its runs + the run index of every row: META aux_off = first run of the vector
 *                     among the chunk's runs, aux_cnt = its runs; DATA = [128 lane bases][128*W deltas] of the u16 run index
 *                     (as numeric RLE); AUX0 = bytes of all run strings; AUX1 = u32 end[aux_count]; aux_count = runs of the chunk
 *   CROSS_RLE_STR:    runs over the whole chunk (padding rows of the last vector extend the last run): AUX0 = bytes of the run
 *                     strings; AUX1 = u32 end[aux_count] | u32 length[aux_count] (lengths sum to n_vec*1024)
 *   FREQUENCY_STR:    per vector one frequent string and the rows that differ: META aux_off = index of the vector's frequent
 *                     string among the chunk's value strings, aux_cnt = n exceptions, whose strings are values aux_off+1 ..
 *                     aux_off+n in position order; DATA blob = u16 position[n] ascending, padded to 16 B; AUX0 = bytes of all
 *                     value strings; AUX1 = u32 end[aux_count]; aux_count = value strings of the chunk (n_vec + all exceptions)
 * const_val = number of DECODED string bytes of the chunk (its share of the string heap); the per-string decoded end
 * offsets make every string's heap position known without decoding its predecessors (entry points, like PointTo). */
typedef struct fls_fsst_table { /* fsst_decoder_t of duckdb/third_party/fsst/fsst.h:70-75 without its header bytes */
	uint64_t symbol[256]; /* symbol[c]: up to 8 bytes, little endian; c = 255 is the escape code and unused */
	uint8_t len[256];     /* byte length of symbol[c], 1..8 */
} fls_fsst_table;
#define FLS_FSST_ESC 255u
/* FSST12 (reference fsst12_kernel.hpp:14-57, fsst12_dict_kernel.hpp; decoder of cwida/fsst's fsst12.h, which lives in the
 * absent FastLanes library): 12-bit codes, two per 3 bytes, little endian: code0 = b0 | (b1 & 15) << 8, code1 = b1 >> 4 |
 * b2 << 4; a string with an odd number of codes ends with one code in 2 bytes.  Codes < 256 stand for the byte itself,
 * codes >= 256 for symbol[code] (2..8 bytes, zero padded).  Segment layout as FSST / FSST_DICT with this table in front. */
typedef struct fls_fsst12_table {
	uint64_t symbol[4096];
	uint8_t len[4096];
} fls_fsst12_table;
#define FLS_FSST_TABLE_BYTES(e) (FLS_ENC_IS_FSST12(e) ? sizeof(fls_fsst12_table) : sizeof(fls_fsst_table))

#define FLS_PAD16(x) (((x) + 15u) & ~15u)
#define FLS_EXC_TAIL_BYTES(n, elem_bytes) ((n) ? FLS_PAD16((uint32_t)(n) * (elem_bytes)) + FLS_PAD16((uint32_t)(n) * 2u) : 0u)

#ifdef __cplusplus
}
#endif
#endif /* FLS_FORMAT_H */
