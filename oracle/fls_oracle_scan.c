/*
 * fls_oracle_scan.c — materializer, column decode, reference-shaped scan and aggregates of the CPU oracle.
 * TEST INFRASTRUCTURE ONLY (see fls_oracle.h).
 */
#include "fls_oracle.h"

#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* same numbering as flsgpu_out_type */
enum { OUT_NATIVE = 0, OUT_I8, OUT_I16, OUT_I32, OUT_I64, OUT_U8, OUT_U16, OUT_U32, OUT_U64, OUT_F32, OUT_F64 };

static int out_bits(int out_type, const fls_column_desc* c) {
	switch (out_type) {
	case OUT_NATIVE: return c->phys_bits;
	case OUT_I8: case OUT_U8: return 8;
	case OUT_I16: case OUT_U16: return 16;
	case OUT_I32: case OUT_U32: case OUT_F32: return 32;
	default: return 64;
	}
}

static inline uint64_t load_src(const void* p, unsigned T, unsigned i, int is_signed) {
	switch (T) {
	case 8: return is_signed ? (uint64_t)(int64_t)((const int8_t*)p)[i] : ((const uint8_t*)p)[i];
	case 16: return is_signed ? (uint64_t)(int64_t)((const int16_t*)p)[i] : ((const uint16_t*)p)[i];
	case 32: return is_signed ? (uint64_t)(int64_t)((const int32_t*)p)[i] : ((const uint32_t*)p)[i];
	default: return ((const uint64_t*)p)[i];
	}
}

/*
 * KernelTraits<Op>::Decode -> NumericHelper::{CopyVector,UntransposeBlock} (fls_type_resolver.hpp:120-165):
 * same physical type -> fastlanes::copy; otherwise CastNumeric's plain static_cast loop (:22-33), with integer
 * columns of signed src_type viewed as signed first (unffor_kernel.hpp:16-25) so that widening sign-extends.
 * A transposed scratch goes through untranspose_i, via a 1024-element temporary when a cast follows (:206-222).
 */
int fls_oracle_materialize(const fls_oracle_table* t, uint32_t col, const void* scratch, int is_transposed, int out_type,
                           void* out) {
	const fls_column_desc* c = &t->columns[col];
	const unsigned T = c->phys_bits;
	const int ob = out_bits(out_type, c);
	const int out_float = out_type == OUT_F32 || out_type == OUT_F64 || (out_type == OUT_NATIVE && c->is_float);
	uint64_t tmp[1024];
	const void* src = scratch;
	if (is_transposed) {
		if ((unsigned)ob == T && out_float == c->is_float) {
			fls_oracle_untranspose(scratch, T, out);
			return 0;
		}
		fls_oracle_untranspose(scratch, T, tmp);
		src = tmp;
	}
	if ((unsigned)ob == T && out_float == c->is_float) {
		memcpy(out, src, 1024u * T / 8);
		return 0;
	}
	if ((unsigned)ob < T) {
		return -5; /* narrowing is never requested by the reader: the bound type is the promoted one */
	}
	if (c->is_float) {
		if (T == 32 && out_type == OUT_F64) {
			for (unsigned i = 0; i < 1024; ++i) {
				((double*)out)[i] = (double)((const float*)src)[i];
			}
			return 0;
		}
		return -5;
	}
	if (out_float) {
		for (unsigned i = 0; i < 1024; ++i) {
			const uint64_t v = load_src(src, T, i, c->is_signed);
			const double d = c->is_signed ? (double)(int64_t)v : (double)v;
			if (ob == 32) {
				((float*)out)[i] = c->is_signed ? (float)(int64_t)v : (float)v;
			} else {
				((double*)out)[i] = d;
			}
		}
		return 0;
	}
	for (unsigned i = 0; i < 1024; ++i) {
		const uint64_t v = load_src(src, T, i, c->is_signed);
		switch (ob) {
		case 16: ((uint16_t*)out)[i] = (uint16_t)v; break;
		case 32: ((uint32_t*)out)[i] = (uint32_t)v; break;
		default: ((uint64_t*)out)[i] = v; break;
		}
	}
	return 0;
}

/* ---- threaded drivers ---------------------------------------------------------------------------------- */
typedef struct job {
	const fls_oracle_table* t;
	const uint32_t* cols;
	uint32_t n_cols, rg_begin, n_rg;
	int out_type;
	uint8_t* out; /* decode_column only */
	int mode;     /* 0: decode column into out; 1: reference-shaped scan */
	volatile uint32_t* next;
	int rc;
	uint64_t bytes, checksum;
} job;

static uint64_t fold(const uint64_t* p, size_t n_words, uint64_t h) {
	for (size_t i = 0; i < n_words; ++i) {
		h = (h ^ p[i]) * 0x100000001B3ull;
	}
	return h;
}

static void* worker(void* arg) {
	job* j = (job*)arg;
	const fls_oracle_table* t = j->t;
	const uint32_t vpr = t->header.vec_per_rowgroup;
	uint64_t scratch[1024];
	/* one 2048-row DuckDB chunk buffer per projected column (vector_size.hpp:16) */
	/* (string columns: 16-byte string_t or 4-byte selection entries; the slot is sized for string_t) */
	uint64_t* chunkbuf = j->mode ? (uint64_t*)malloc((size_t)j->n_cols * 2048 * 16) : NULL;
	/* string columns: per-column dictionary (keys + bytes, rebuilt per row group like dictionary_kernel.hpp:60-78 Prepare)
	 * and the per-chunk string heap of flat vectors (the reference allocates per string) */
	uint8_t** str_keys = j->mode ? (uint8_t**)calloc(j->n_cols, sizeof(uint8_t*)) : NULL;
	uint8_t** str_heap = j->mode ? (uint8_t**)calloc(j->n_cols, sizeof(uint8_t*)) : NULL;
	uint64_t* str_cap = j->mode ? (uint64_t*)calloc((size_t)j->n_cols * 2, sizeof(uint64_t)) : NULL;
	for (;;) {
		/* TryAssignNextRowGroup: row groups are handed out one at a time (fls_reader.cpp:337-355) */
		const uint32_t r = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED);
		if (r >= j->n_rg) {
			break;
		}
		const uint32_t rg = j->rg_begin + r;
		const uint32_t n_vec = t->rowgroups[rg].n_vec;
		if (!j->mode) {
			const fls_column_desc* c = &t->columns[j->cols[0]];
			const size_t esz = (size_t)out_bits(j->out_type, c) / 8;
			for (uint32_t v = 0; v < n_vec; ++v) {
				int tr = 0;
				int rc = fls_oracle_decode_vector(t, rg, j->cols[0], v, scratch, &tr);
				if (!rc) {
					rc = fls_oracle_materialize(t, j->cols[0], scratch, tr, j->out_type,
					                            j->out + ((size_t)r * vpr + v) * 1024 * esz);
				}
				if (rc) {
					j->rc = rc;
					return NULL;
				}
			}
		} else {
			/* Scan: GetScanBatch takes <= 2 vectors per call (fls_reader.cpp:422-434), DecodePhysicalColumns visits
			 * every projected column per vector (:436-472) with Pass::First / Pass::Second destinations */
			/* InitializeColumnDecoders -> Prepare (fls_reader.cpp:382-420): string dictionaries are built once per row group */
			for (uint32_t ci = 0; ci < j->n_cols; ++ci) {
				const fls_column_desc* c = &t->columns[j->cols[ci]];
				if (c->phys_bits != 128) {
					continue;
				}
				const fls_chunk_desc* ch = &t->chunks[(size_t)rg * t->header.n_columns + j->cols[ci]];
				const int is_dict = FLS_ENC_IS_STR_DICT(ch->encoding);
				const uint64_t need_keys = is_dict ? (uint64_t)ch->aux_count * 16 + 16 : 0;
				uint64_t need_heap = (is_dict ? ch->const_val : ch->encoding == FLS_ENC_STR_CONSTANT ? (uint64_t)ch->aux_count : ch->const_val) + 64;
				if (FLS_ENC_IS_STR_VALUES(ch->encoding)) { /* AddString per row: a vector's heap holds up to 1024 copies of the longest value */
					const uint32_t* ends = (const uint32_t*)(t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX1]);
					uint32_t longest = 0;
					for (uint32_t k = 0; k < ch->aux_count; ++k) {
						const uint32_t len = ends[k] - (k ? ends[k - 1] : 0);
						longest = len > longest ? len : longest;
					}
					need_heap = (uint64_t)longest * FLS_VEC_SZ + 64;
				}
				if (str_cap[2 * ci] < need_keys) {
					free(str_keys[ci]);
					str_keys[ci] = (uint8_t*)malloc(need_keys);
					str_cap[2 * ci] = need_keys;
				}
				if (str_cap[2 * ci + 1] < need_heap) {
					free(str_heap[ci]);
					str_heap[ci] = (uint8_t*)malloc(need_heap);
					str_cap[2 * ci + 1] = need_heap;
				}
				if (is_dict) {
					const int rc = fls_oracle_decode_string_dict(t, j->cols[ci], rg, str_keys[ci], str_heap[ci], (uint64_t)(uintptr_t)str_heap[ci]);
					if (rc) {
						j->rc = rc;
						goto done;
					}
					j->checksum = fold((const uint64_t*)str_keys[ci], (size_t)ch->aux_count * 2, j->checksum);
				}
			}
			for (uint32_t v0 = 0; v0 < n_vec; v0 += 2) {
				const uint32_t nv = n_vec - v0 < 2 ? n_vec - v0 : 2;
				for (uint32_t b = 0; b < nv; ++b) {
					for (uint32_t ci = 0; ci < j->n_cols; ++ci) {
						const fls_column_desc* c = &t->columns[j->cols[ci]];
						if (c->phys_bits == 128) { /* string column: selection vector over the dictionary, or flat string_t */
							const fls_chunk_desc* ch = &t->chunks[(size_t)rg * t->header.n_columns + j->cols[ci]];
							int rc;
							if (FLS_ENC_IS_STR_DICT(ch->encoding)) {
								uint32_t* sel = (uint32_t*)(chunkbuf + (size_t)ci * 4096) + (size_t)b * 1024;
								rc = fls_oracle_decode_codes(t, rg, j->cols[ci], v0 + b, sel);
								j->checksum = fold((const uint64_t*)sel, 512, j->checksum);
								j->bytes += 4096;
							} else {
								uint8_t* dst = (uint8_t*)(chunkbuf + (size_t)ci * 4096) + (size_t)b * 1024 * 16;
								uint64_t used = 0;
								rc = fls_oracle_decode_string_vector(t, rg, j->cols[ci], v0 + b, dst, str_heap[ci],
								                                     (uint64_t)(uintptr_t)str_heap[ci], &used);
								j->checksum = fold((const uint64_t*)str_heap[ci], used / 8, j->checksum);
								j->bytes += 16384 + used;
							}
							if (rc) {
								j->rc = rc;
								goto done;
							}
							continue;
						}
						const size_t esz = (size_t)c->phys_bits / 8;
						uint8_t* dst = (uint8_t*)(chunkbuf + (size_t)ci * 4096) + (size_t)b * 1024 * esz;
						int tr = 0;
						int rc = fls_oracle_decode_vector(t, rg, j->cols[ci], v0 + b, scratch, &tr);
						if (!rc) {
							rc = fls_oracle_materialize(t, j->cols[ci], scratch, tr, OUT_NATIVE, dst);
						}
						if (rc) {
							j->rc = rc;
							goto done;
						}
					}
				}
				for (uint32_t ci = 0; ci < j->n_cols; ++ci) {
					if (t->columns[j->cols[ci]].phys_bits == 128) {
						continue; /* counted above */
					}
					const size_t esz = (size_t)t->columns[j->cols[ci]].phys_bits / 8;
					const size_t nbytes = (size_t)nv * 1024 * esz;
					j->checksum = fold(chunkbuf + (size_t)ci * 4096, nbytes / 8, j->checksum);
					j->bytes += nbytes;
				}
			}
		}
	}
done:
	free(chunkbuf);
	for (uint32_t ci = 0; str_keys && ci < j->n_cols; ++ci) {
		free(str_keys[ci]);
		free(str_heap[ci]);
	}
	free(str_keys);
	free(str_heap);
	free(str_cap);
	return NULL;
}

static int run_jobs(job* proto, int n_threads) {
	if (n_threads < 1) {
		n_threads = 1;
	}
	volatile uint32_t next = 0;
	job* jobs = (job*)calloc((size_t)n_threads, sizeof(job));
	pthread_t* th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
	for (int i = 0; i < n_threads; ++i) {
		jobs[i] = *proto;
		jobs[i].next = &next;
		jobs[i].checksum = 0xcbf29ce484222325ull + (uint64_t)i;
		if (n_threads > 1) {
			pthread_create(&th[i], NULL, worker, &jobs[i]);
		}
	}
	if (n_threads == 1) {
		worker(&jobs[0]);
	}
	int rc = 0;
	proto->bytes = 0;
	proto->checksum = 0;
	for (int i = 0; i < n_threads; ++i) {
		if (n_threads > 1) {
			pthread_join(th[i], NULL);
		}
		if (jobs[i].rc) {
			rc = jobs[i].rc;
		}
		proto->bytes += jobs[i].bytes;
		proto->checksum ^= jobs[i].checksum;
	}
	free(jobs);
	free(th);
	return rc;
}

int fls_oracle_decode_column(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, int out_type,
                             void* out, int n_threads) {
	if (col >= t->header.n_columns || rg_begin + n_rg > t->header.n_rowgroups) {
		return -1;
	}
	job j;
	memset(&j, 0, sizeof(j));
	j.t = t;
	j.cols = &col;
	j.n_cols = 1;
	j.rg_begin = rg_begin;
	j.n_rg = n_rg;
	j.out_type = out_type;
	j.out = (uint8_t*)out;
	j.mode = 0;
	return run_jobs(&j, n_threads);
}

int fls_oracle_scan(const fls_oracle_table* t, const uint32_t* cols, uint32_t n_cols, uint32_t rg_begin, uint32_t n_rg,
                    int n_threads, uint64_t* decoded_bytes, uint64_t* checksum) {
	if (rg_begin + n_rg > t->header.n_rowgroups) {
		return -1;
	}
	for (uint32_t i = 0; i < n_cols; ++i) {
		if (cols[i] >= t->header.n_columns) {
			return -1;
		}
	}
	job j;
	memset(&j, 0, sizeof(j));
	j.t = t;
	j.cols = cols;
	j.n_cols = n_cols;
	j.rg_begin = rg_begin;
	j.n_rg = n_rg;
	j.mode = 1;
	const int rc = run_jobs(&j, n_threads);
	if (decoded_bytes) {
		*decoded_bytes = j.bytes;
	}
	if (checksum) {
		*checksum = j.checksum;
	}
	return rc;
}

/* SUM as 128-bit two's complement (DuckDB SUM(BIGINT)->HUGEINT, sum.cpp:161-210), COUNT, MIN, MAX over the valid
 * tuples of the range. */
int fls_oracle_aggregate(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* sum_lo,
                         uint64_t* sum_hi, uint64_t* count, int64_t* min_val, int64_t* max_val, double* fsum) {
	if (col >= t->header.n_columns || rg_begin + n_rg > t->header.n_rowgroups) {
		return -1;
	}
	const fls_column_desc* c = &t->columns[col];
	const unsigned T = c->phys_bits;
	unsigned __int128 acc = 0;
	uint64_t n = 0;
	int first = 1;
	int64_t mn = 0, mx = 0;
	double fs = 0.0;
	uint64_t scratch[1024], nat[1024];
	for (uint32_t r = 0; r < n_rg; ++r) {
		const uint32_t rg = rg_begin + r;
		uint32_t left = t->rowgroups[rg].n_tuples;
		for (uint32_t v = 0; v < t->rowgroups[rg].n_vec; ++v) {
			int tr = 0;
			int rc = fls_oracle_decode_vector(t, rg, col, v, scratch, &tr);
			if (!rc) {
				rc = fls_oracle_materialize(t, col, scratch, tr, OUT_NATIVE, nat);
			}
			if (rc) {
				return rc;
			}
			const uint32_t valid = left < 1024 ? left : 1024;
			left -= valid;
			for (uint32_t i = 0; i < valid; ++i) {
				if (c->is_float) {
					double x = T == 32 ? (double)((const float*)nat)[i] : ((const double*)nat)[i];
					fs += x;
					continue;
				}
				const uint64_t u = load_src(nat, T, i, c->is_signed);
				if (c->is_signed) {
					acc += (unsigned __int128)(__int128)(int64_t)u;
					if (first || (int64_t)u < mn) mn = (int64_t)u;
					if (first || (int64_t)u > mx) mx = (int64_t)u;
				} else {
					acc += u;
					if (first || u < (uint64_t)mn) mn = (int64_t)u;
					if (first || u > (uint64_t)mx) mx = (int64_t)u;
				}
				first = 0;
			}
			n += valid;
		}
	}
	*sum_lo = (uint64_t)acc;
	*sum_hi = (uint64_t)(acc >> 64);
	*count = n;
	*min_val = mn;
	*max_val = mx;
	*fsum = fs;
	return 0;
}

/* ---- filter pushdown ------------------------------------------------------------------------------------------------
 * FilterExecutor::Apply (reference src/reader/filter_executor.cpp:19-60) hands every filter column of a decoded chunk to
 * DuckDB's ColumnSegment::FilterSelection; for a ConstantFilter that is a comparison of each value with the constant in
 * the column's type (duckdb/src/storage/table/column_segment.cpp, templated_filter_operation), floating point with
 * DuckDB's total order (NaN == NaN, NaN greater than everything: duckdb/common/operator/comparison_operators.hpp).
 * The selection is returned as a bit mask in decode_column's row layout: bit i of word w of vector v. */
static int cmp_values(const fls_column_desc* c, uint64_t a, uint64_t b) {
	if (c->is_float) {
		double x, y;
		if (c->phys_bits == 32) {
			float fx, fy;
			const uint32_t ua = (uint32_t)a, ub = (uint32_t)b;
			memcpy(&fx, &ua, 4);
			memcpy(&fy, &ub, 4);
			x = fx;
			y = fy;
		} else {
			memcpy(&x, &a, 8);
			memcpy(&y, &b, 8);
		}
		const int xn = x != x, yn = y != y;
		if (xn || yn) {
			return xn && yn ? 0 : xn ? 1 : -1;
		}
		return x < y ? -1 : x > y ? 1 : 0;
	}
	if (c->is_signed) {
		return (int64_t)a < (int64_t)b ? -1 : (int64_t)a > (int64_t)b ? 1 : 0;
	}
	return a < b ? -1 : a > b ? 1 : 0;
}

static int passes(int op, int cmp) {
	switch (op) {
	case 0: return cmp == 0;
	case 1: return cmp != 0;
	case 2: return cmp < 0;
	case 3: return cmp <= 0;
	case 4: return cmp > 0;
	default: return cmp >= 0;
	}
}

/* VARCHAR / BLOB constants: DuckDB compares string_t bytewise (memcmp over the common length), the shorter string first
 * on a tie (duckdb/src/include/duckdb/common/types/string_type.hpp:155-217, StringComparisonOperators: GreaterThan is
 * memcmp(left, right, min_len) > 0 || (== 0 && left_len > right_len); Equals compares length + bytes).  The filter column
 * is decoded the way the scan decodes it (flat string_t rows of one row group) and every valid row is compared. */
typedef struct ofilter_str {
	uint32_t len;
	uint8_t prefix[4];
	union {
		uint8_t inlined[8];
		uint64_t ptr;
	} v;
} ofilter_str;

static int cmp_string(const ofilter_str* s, const uint8_t* k, uint64_t klen) {
	const uint8_t* p = s->len <= 12 ? s->prefix : (const uint8_t*)(uintptr_t)s->v.ptr;
	const uint64_t m = s->len < klen ? s->len : klen;
	const int c = m ? memcmp(p, k, m) : 0;
	if (c) {
		return c < 0 ? -1 : 1;
	}
	return s->len < klen ? -1 : s->len > klen ? 1 : 0;
}

static int filter_string_rowgroup(const fls_oracle_table* t, const fls_oracle_filter* f, uint32_t rg, uint32_t* mask_rg) {
	const uint32_t vpr = t->header.vec_per_rowgroup;
	uint64_t heap_bytes = 0;
	if (fls_oracle_string_heap_bytes(t, f->col, rg, 1, &heap_bytes)) {
		return -1;
	}
	ofilter_str* rows = (ofilter_str*)malloc((size_t)vpr * 1024 * sizeof(ofilter_str));
	uint8_t* heap = (uint8_t*)malloc(heap_bytes + 16);
	int rc = rows && heap ? fls_oracle_decode_strings(t, f->col, rg, 1, rows, heap, (uint64_t)(uintptr_t)heap) : -1;
	if (!rc) {
		const uint32_t n = t->rowgroups[rg].n_tuples;
		for (uint32_t i = 0; i < n; ++i) {
			if (!passes(f->op, cmp_string(&rows[i], f->str_data, f->str_len))) {
				mask_rg[i >> 5] &= ~(1u << (i & 31));
			}
		}
	}
	free(rows);
	free(heap);
	return rc;
}

int fls_oracle_filter_mask(const fls_oracle_table* t, const fls_oracle_filter* filters, uint32_t n_filters, uint32_t rg_begin,
                           uint32_t n_rg, uint32_t* mask) {
	if (rg_begin + n_rg > t->header.n_rowgroups) {
		return -1;
	}
	const uint32_t vpr = t->header.vec_per_rowgroup;
	memset(mask, 0, (size_t)n_rg * vpr * 128);
	uint64_t scratch[1024], nat[1024];
	for (uint32_t r = 0; r < n_rg; ++r) {
		const uint32_t rg = rg_begin + r;
		uint32_t left = t->rowgroups[rg].n_tuples;
		for (uint32_t v = 0; v < t->rowgroups[rg].n_vec; ++v) {
			const uint32_t valid = left < 1024 ? left : 1024;
			left -= valid;
			uint32_t* mv = mask + ((size_t)r * vpr + v) * 32;
			for (uint32_t i = 0; i < valid; ++i) {
				mv[i >> 5] |= 1u << (i & 31);
			}
			for (uint32_t f = 0; f < n_filters; ++f) {
				if (filters[f].col >= t->header.n_columns) {
					return -1;
				}
				if (t->columns[filters[f].col].phys_bits == 128) {
					continue; /* string filters: one pass per row group, below */
				}
				const fls_column_desc* c = &t->columns[filters[f].col];
				int tr = 0;
				int rc = fls_oracle_decode_vector(t, rg, filters[f].col, v, scratch, &tr);
				if (!rc) {
					rc = fls_oracle_materialize(t, filters[f].col, scratch, tr, OUT_NATIVE, nat);
				}
				if (rc) {
					return rc;
				}
				for (uint32_t i = 0; i < valid; ++i) {
					const uint64_t val = c->is_float ? (c->phys_bits == 32 ? ((const uint32_t*)nat)[i] : nat[i])
					                                 : load_src(nat, c->phys_bits, i, c->is_signed);
					if (!passes(filters[f].op, cmp_values(c, val, filters[f].constant))) {
						mv[i >> 5] &= ~(1u << (i & 31));
					}
				}
			}
		}
		for (uint32_t f = 0; f < n_filters; ++f) {
			if (t->columns[filters[f].col].phys_bits == 128) {
				if ((!filters[f].str_data && filters[f].str_len) || filter_string_rowgroup(t, &filters[f], rg, mask + (size_t)r * vpr * 32)) {
					return -1;
				}
			}
		}
	}
	return 0;
}

/* RowGroupFilter::RowGroupMaySatisfyFilters (reference src/reader/row_group_filter.cpp:96-199): only <, <=, >, >= against
 * the row group's min / max statistic prune. */
int fls_oracle_rowgroup_may_satisfy(const fls_oracle_table* t, const fls_oracle_filter* filters, uint32_t n_filters, uint32_t rg) {
	for (uint32_t f = 0; f < n_filters; ++f) {
		const fls_column_desc* c = &t->columns[filters[f].col];
		const fls_chunk_desc* ch = &t->chunks[(size_t)rg * t->header.n_columns + filters[f].col];
		if (!ch->has_stats || c->is_float || c->phys_bits == 128) {
			continue; /* no zone maps on floating point and string chunks */
		}
		const int cmin = cmp_values(c, ch->min_val, filters[f].constant), cmax = cmp_values(c, ch->max_val, filters[f].constant);
		switch (filters[f].op) {
		case 4: if (cmax <= 0) return 0; break; /* > : skip when max <= constant */
		case 5: if (cmax < 0) return 0; break;
		case 2: if (cmin >= 0) return 0; break;
		case 3: if (cmin > 0) return 0; break;
		default: break;
		}
	}
	return 1;
}

/* SUM (128-bit) / COUNT of the rows of `col` whose mask bit is set. */
int fls_oracle_aggregate_masked(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const uint32_t* mask,
                                uint64_t* sum_lo, uint64_t* sum_hi, uint64_t* count, double* fsum) {
	if (col >= t->header.n_columns || rg_begin + n_rg > t->header.n_rowgroups) {
		return -1;
	}
	const fls_column_desc* c = &t->columns[col];
	const unsigned T = c->phys_bits;
	const uint32_t vpr = t->header.vec_per_rowgroup;
	unsigned __int128 acc = 0;
	uint64_t n = 0;
	double fs = 0.0;
	uint64_t scratch[1024], nat[1024];
	for (uint32_t r = 0; r < n_rg; ++r) {
		const uint32_t rg = rg_begin + r;
		for (uint32_t v = 0; v < t->rowgroups[rg].n_vec; ++v) {
			const uint32_t* mv = mask + ((size_t)r * vpr + v) * 32;
			int tr = 0;
			int rc = fls_oracle_decode_vector(t, rg, col, v, scratch, &tr);
			if (!rc) {
				rc = fls_oracle_materialize(t, col, scratch, tr, OUT_NATIVE, nat);
			}
			if (rc) {
				return rc;
			}
			for (uint32_t i = 0; i < 1024; ++i) {
				if (!((mv[i >> 5] >> (i & 31)) & 1u)) {
					continue;
				}
				++n;
				if (c->is_float) {
					fs += T == 32 ? (double)((const float*)nat)[i] : ((const double*)nat)[i];
					continue;
				}
				const uint64_t u = load_src(nat, T, i, c->is_signed);
				acc += c->is_signed ? (unsigned __int128)(__int128)(int64_t)u : (unsigned __int128)u;
			}
		}
	}
	*sum_lo = (uint64_t)acc;
	*sum_hi = (uint64_t)(acc >> 64);
	*count = n;
	*fsum = fs;
	return 0;
}
