// flsgpu_kernels_str.cu — string column kernels (sm_100a): FSST decode, string heap fill, string_t materialisation,
// dictionary codes (selection vectors).
//
// Replaces, for VARCHAR / BLOB columns, the reference materializer's string kernels
// (reference src/include/reader/materializer/kernels/): uncompressed_kernel.hpp:26-38, constant_kernel.hpp:40-49,
// dictionary_kernel.hpp:52-102, fsst_kernel.hpp:19-57, fsst_dict_kernel.hpp:20-80 — which build 1024 string_t per vector
// with one heap allocation + one scalar fsst_decompress (duckdb/third_party/fsst/fsst.h:162-230) per string.
//
// Here a string column of a row-group range is decoded by two launches:
//   1. fls_str_heap_kernel: every chunk's decoded bytes go to its slice of ONE string heap (the chunk's slice starts at the
//      16-byte padded prefix sum of the chunks before it).  Plain encodings are a 16-byte-granule copy; FSST chunks are
//      decoded in parallel: the per-string entry points (compressed end, decoded end) let every lane start at its own
//      string, a lane decodes a run of consecutive strings = one contiguous sub-stream, assembles the output in a 64-bit
//      register and writes aligned 8-byte words (only the first/last partial word of a lane's run is written bytewise,
//      because it is shared with the neighbouring lane).  The 255-symbol table sits in shared memory.
//   2. fls_str_rows_kernel: one warp per 1024-row vector builds DuckDB string_t (string_type.hpp:22-60,236-247): length,
//      then 12 inlined bytes (zero padded) or 4 prefix bytes + pointer.  Pointers are heap_base_addr + offset, i.e. already
//      valid for wherever the caller is going to place the heap (its own host buffer after the D2H copy, or the device heap).
//      Dictionary encodings unpack their codes (same FFOR blobs as numeric DICT) through a per-warp shared-memory tile.
// fls_str_codes_kernel emits the u32 selection vector of dictionary-encoded chunks (dictionary_kernel.hpp:80-101: the
// reference hands DuckDB a DICTIONARY_VECTOR = keys + selection vector instead of flat strings).
#include <cuda_runtime.h>

#include "fls_launch.hpp"
#include "fls_ptx.cuh"
#include "fls_unpack.cuh"

using namespace flsk;

namespace {

constexpr int kWarps = 8;

__device__ __forceinline__ bool is_fsst(uint32_t enc) { // FSST or FSST12, per row or dictionary
	return FLS_ENC_IS_FSST_ANY(enc);
}

// Output side of a lane's decode: symbols are ORed into a 64-bit register at the current byte offset and flushed as aligned
// 8-byte words; the first / last partial word of the lane's run is written bytewise (the neighbouring lanes own the rest).
struct LaneOut {
	unsigned long long* ow;
	unsigned long long acc;
	uint32_t fill, head; // head: low bytes of the first word that belong to the previous lane's strings
	__device__ __forceinline__ LaneOut(uint8_t* out, uint32_t dbeg) {
		const uint32_t mis = dbeg & 7u;
		ow = reinterpret_cast<unsigned long long*>(out + (dbeg - mis));
		acc = 0;
		fill = head = mis;
	}
	__device__ __forceinline__ void append(unsigned long long sym, uint32_t len) { // sym zero padded beyond len (1..8)
		acc |= sym << (8u * fill);
		uint32_t nf = fill + len;
		if (nf >= 8u) {
			if (head) {
				uint8_t* ob = reinterpret_cast<uint8_t*>(ow);
				for (uint32_t k = head; k < 8u; ++k) {
					ob[k] = (uint8_t)(acc >> (8u * k));
				}
				head = 0;
			} else {
				*ow = acc;
			}
			++ow;
			nf -= 8u;
			acc = nf ? (sym >> (8u * (len - nf))) : 0ull;
		}
		fill = nf;
	}
	__device__ __forceinline__ void finish() {
		uint8_t* ob = reinterpret_cast<uint8_t*>(ow);
		for (uint32_t k = head; k < fill; ++k) {
			ob[k] = (uint8_t)(acc >> (8u * k));
		}
	}
};

// FSST12 (fls_format.h fls_fsst12_table; reference fsst12_kernel.hpp:53): one lane decodes its strings [first, last] one
// after the other - a string's stream starts on a byte boundary, packs two 12-bit codes per 3 bytes and ends with a
// 2-byte single code when its code count is odd.  The 36 KiB table stays in global memory (read-only path, L1 resident).
template <bool GLOBAL_IN>
__device__ __forceinline__ void fsst12_decode_lane(const uint8_t* __restrict__ in, const uint32_t* comp_end, uint32_t first, uint32_t last,
                                                   uint32_t cbeg, uint8_t* __restrict__ out, uint32_t dbeg, const fls_fsst12_table* tab) {
	LaneOut o(out, dbeg);
	uint32_t p = cbeg;
	for (uint32_t s = first; s <= last; ++s) {
		const uint32_t ce = __ldg(comp_end + s);
		while (p + 2u <= ce) {
			const uint32_t b0 = GLOBAL_IN ? __ldg(in + p) : in[p], b1 = GLOBAL_IN ? __ldg(in + p + 1) : in[p + 1];
			const uint32_t c0 = b0 | ((b1 & 15u) << 8);
			o.append(__ldg(reinterpret_cast<const unsigned long long*>(&tab->symbol[c0])), __ldg(&tab->len[c0]));
			if (p + 3u <= ce) {
				const uint32_t b2 = GLOBAL_IN ? __ldg(in + p + 2) : in[p + 2];
				const uint32_t c1 = (b1 >> 4) | (b2 << 4);
				o.append(__ldg(reinterpret_cast<const unsigned long long*>(&tab->symbol[c1])), __ldg(&tab->len[c1]));
				p += 3u;
			} else {
				p += 2u;
			}
		}
		p = ce;
	}
	o.finish();
}

// One lane decodes the code stream [cbeg, cend) (relative to `in`) to out[dbeg ...).  `out` is 8-byte aligned.
// fsst.h:219-230 per code; output assembled in a 64-bit register, flushed as aligned 64-bit words.
// GLOBAL_IN: the codes are read straight from global memory (read-only path); otherwise from the warp's staging buffer.
// tab: shared-memory address of the 256 x 16-byte decoder table {symbol (zero padded), length in BITS}: one 128-bit load
// per code.  The loop is issue bound, so it is written for instruction count: the next code byte is loaded while the
// current one is looked up, the fill level is kept in bits, the carry-over after a flush is two shifts (no selects), and
// the table is addressed from a register (no per-iteration recomputation of shared-memory bases).
// position in a code stream: a pointer (global memory) or a 32-bit shared-memory address (staging buffer)
template <bool GLOBAL_IN>
struct FsstPos;
template <>
struct FsstPos<true> {
	const uint8_t* p;
	__device__ __forceinline__ FsstPos(const uint8_t* q) : p(q) {}
	__device__ __forceinline__ uint32_t at(uint32_t k) const { return __ldg(p + k); }
};
template <>
struct FsstPos<false> {
	uint32_t p;
	__device__ __forceinline__ FsstPos(const uint8_t* q) : p((uint32_t)__cvta_generic_to_shared(q)) {}
	__device__ __forceinline__ uint32_t at(uint32_t k) const {
		uint32_t v;
		asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(p + k));
		return v;
	}
};

template <bool GLOBAL_IN>
__device__ __forceinline__ void fsst_decode_lane(const uint8_t* __restrict__ in, uint32_t cbeg, uint32_t cend, uint8_t* __restrict__ out,
                                                 uint32_t dbeg, uint32_t tab) {
	const uint32_t mis = dbeg & 7u;
	unsigned long long* ow = reinterpret_cast<unsigned long long*>(out + (dbeg - mis));
	unsigned long long acc = 0;
	uint32_t fill8 = 8u * mis, head = mis; // head: low bytes of the first word that belong to the previous lane's strings
	if (cbeg < cend) {
		FsstPos<GLOBAL_IN> pos(in + cbeg);
		auto& p = pos.p;
		const auto pe = p + (cend - cbeg);
		asm volatile("mov.u32 %0, %0;" : "+r"(tab)); // keep the table address in a register (else it is recomputed per code)
		// a full word leaves.  Staged tiles: always as one 8-byte store - the low bytes of a lane's first word belong to the
		// previous lane, which writes them AFTER the warp has synchronised (below), over the zeros stored here.  Straight to
		// the heap (GLOBAL_IN) the first word may be shared with another warp's tile, so its bytes are stored one by one.
		auto flush = [&](unsigned long long sym, uint32_t len8, uint32_t nf8) {
			if (GLOBAL_IN && head) {
				uint8_t* ob = reinterpret_cast<uint8_t*>(ow);
				for (uint32_t k = head; k < 8u; ++k) {
					ob[k] = (uint8_t)(acc >> (8u * k));
				}
			} else {
				*ow = acc;
			}
			head = 0;
			++ow;
			nf8 -= 64u;
			// what did not fit: sym >> (bits that went into the flushed word) - 8..64 of them, 64 when an 8-byte symbol
			// filled an empty word, hence the shift in two steps
			acc = (sym >> (len8 - nf8 - 1u)) >> 1;
			return nf8;
		};
		uint32_t code = pos.at(0);
		for (;;) {
			// hot loop: table codes only, its two exits (escape, end of stream) are real branches
			while (code != FLS_FSST_ESC) {
				++p;
				// speculative: at the end of the stream this reads the byte after it (inside the staging buffer / the segment's slack)
				const uint32_t nxt = pos.at(0);
				uint32_t lo, hi, len8, pad;
				asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo), "=r"(hi), "=r"(len8), "=r"(pad) : "r"(tab + code * 16u));
				const unsigned long long sym = ((unsigned long long)hi << 32) | lo;
				acc |= sym << fill8;
				fill8 += len8;
				if (fill8 >= 64u) {
					fill8 = flush(sym, len8, fill8);
				}
				if (p >= pe) {
					goto done;
				}
				code = nxt;
			}
			{ // escape: the next byte is a literal
				const unsigned long long lit = pos.at(1);
				p += 2;
				acc |= lit << fill8;
				fill8 += 8u;
				if (fill8 >= 64u) {
					fill8 = flush(lit, 8u, fill8);
				}
				if (p >= pe) {
					goto done;
				}
				code = pos.at(0);
			}
		}
	}
done:
	if constexpr (!GLOBAL_IN) {
		__syncwarp(); // every lane's full words are in the buffer: now the partial ones on top of them
	}
	uint8_t* ob = reinterpret_cast<uint8_t*>(ow);
	for (uint32_t k = head; k < (fill8 >> 3); ++k) {
		ob[k] = (uint8_t)(acc >> (8u * k));
	}
}

// DuckDB string_t of the string at [start, start + len) of a chunk's heap slice (hp = where the slice's byte 0 can be
// read: the device heap, or the warp's staging buffer shifted so that slice offsets index it; haddr = the address the
// slice will have where the caller reads the strings).  CG: read through L2 (bytes this warp has just written to global).
template <bool CG>
__device__ __forceinline__ uint4 make_string_t(const uint8_t* hp, uint64_t haddr, uint32_t start, uint32_t len) {
	const uint32_t need = len <= 12u ? len : 4u;
	const uint32_t a = start & ~3u, off = start & 3u, span = off + need;
	const uint32_t* wp = reinterpret_cast<const uint32_t*>(hp + a);
	uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
	if (CG) {
		w0 = span > 0u ? __ldcg(wp) : 0u;
		w1 = span > 4u ? __ldcg(wp + 1) : 0u;
		w2 = span > 8u ? __ldcg(wp + 2) : 0u;
		w3 = span > 12u ? __ldcg(wp + 3) : 0u;
	} else {
		w0 = span > 0u ? wp[0] : 0u;
		w1 = span > 4u ? wp[1] : 0u;
		w2 = span > 8u ? wp[2] : 0u;
		w3 = span > 12u ? wp[3] : 0u;
	}
	const uint32_t sh = off * 8u;
	uint32_t b0 = __funnelshift_r(w0, w1, sh), b1 = __funnelshift_r(w1, w2, sh), b2 = __funnelshift_r(w2, w3, sh);
	b0 &= need >= 4u ? 0xffffffffu : ((1u << (8u * need)) - 1u);
	b1 &= need >= 8u ? 0xffffffffu : need > 4u ? ((1u << (8u * (need - 4u))) - 1u) : 0u;
	b2 &= need >= 12u ? 0xffffffffu : need > 8u ? ((1u << (8u * (need - 8u))) - 1u) : 0u;
	if (len <= 12u) {
		return make_uint4(len, b0, b1, b2);
	}
	const uint64_t ptr = haddr + start;
	return make_uint4(len, b0, (uint32_t)ptr, (uint32_t)(ptr >> 32));
}

// string_t of the tile's strings [s0, s1) (per-row encodings: string index == row index inside the chunk)
template <bool CG>
__device__ __forceinline__ void tile_rows(const uint8_t* hp, uint64_t haddr, const uint32_t* dec_end, uint32_t s0, uint32_t s1, uint32_t d0,
                                          uint32_t lane, uint4* out_chunk) {
	uint32_t carry = d0;
	for (uint32_t s = s0; s < s1; s += 32u) {
		const uint32_t i = s + lane;
		const uint32_t end = i < s1 ? __ldg(dec_end + i) : 0u;
		uint32_t start = __shfl_up_sync(0xffffffffu, end, 1);
		if (lane == 0) {
			start = carry;
		}
		carry = __shfl_sync(0xffffffffu, end, 31);
		if (i < s1) {
			out_chunk[i] = make_string_t<CG>(hp, haddr, start, end - start);
		}
	}
}

// grid: n_rg * groups_per_rg CTAs of 8 warps; CTA (rg, g) takes tiles [8g, 8g + 8) of chunk rg (FSST) or the g-th share
// of its bytes (plain copy).  FSST tile = 32 * lane_strings consecutive strings of the chunk, one contiguous piece of the
// code stream [c0, c1) and of the decoded heap [d0, d1):
//   1. one TMA bulk copy stages the tile's codes in the warp's shared-memory buffer (coalesced, 16-byte granules);
//   2. every lane decodes its lane_strings strings from shared memory into the warp's output buffer (shared memory);
//   3. the buffer goes to the heap with coalesced 16-byte stores (the tile's first / last partial granule bytewise: the
//      neighbouring tiles own the rest of those granules), and - L.fuse_rows - the string_t of the tile's rows are built
//      straight from the buffer, so the heap is written once and never read back.
// Decoding lane-to-global directly (the fallback for tiles that do not fit the buffers) costs 2.8x the DRAM reads and 1.6x
// the writes: 32 lanes x 8-byte stores hit 32 different sectors, partial sector writes make L2 fetch the rest.
__device__ __forceinline__ void str_heap_body(const FlsStrLaunch& L) {
	__shared__ uint4 s_tab[256]; // FSST decoder table: {symbol lo, symbol hi, length in bits, 0}
	__shared__ __align__(8) uint64_t s_bar[2 * kWarps]; // one per (warp, staging buffer)
	extern __shared__ __align__(128) uint8_t s_dyn[];
	const uint32_t rg = blockIdx.x / L.groups_per_rg, g = blockIdx.x - rg * L.groups_per_rg;
	const FlsChunkDev* ch = L.chunks + (size_t)rg * L.chunk_stride;
	const uint32_t enc = ch->encoding;
	const uint64_t hoff = L.heap_prefix[(size_t)rg * L.chunk_stride] - L.heap_prefix0;
	uint8_t* heap = L.heap + hoff;
	if (L.dict_only && !FLS_ENC_IS_STR_DICT(enc)) {
		return;
	}
	if (is_fsst(enc)) {
		const bool twelve = FLS_ENC_IS_FSST12(enc);
		const uint32_t n = ch->aux_count;
		const uint32_t tile_strings = 32u * L.lane_strings;
		if ((uint64_t)g * kWarps * tile_strings >= n) {
			return;
		}
		const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
		{ // symbol table -> shared memory, symbols masked to their length so that OR-ing them into the output word is exact
			// (FSST12: the 4096-entry table is read from global memory, validated zero padded when the table was opened)
			if (!twelve) {
				const fls_fsst_table* tab = reinterpret_cast<const fls_fsst_table*>(ch->aux0);
				const uint32_t c = threadIdx.x;
				const uint32_t len = __ldg(&tab->len[c]);
				const unsigned long long sym = __ldg(reinterpret_cast<const unsigned long long*>(&tab->symbol[c]));
				const unsigned long long m = len >= 8u ? sym : (sym & ((1ull << (8u * len)) - 1ull));
				s_tab[c] = make_uint4((uint32_t)m, (uint32_t)(m >> 32), 8u * len, 0u);
			}
			if (lane == 0) {
				mbar_init(&s_bar[2u * warp], 1);
				mbar_init(&s_bar[2u * warp + 1u], 1);
				fence_mbar_init();
			}
		}
		__syncthreads();
		const uint32_t tab_addr = (uint32_t)__cvta_generic_to_shared(s_tab);
		const uint32_t* comp_end = reinterpret_cast<const uint32_t*>(ch->aux1);
		const uint32_t* dec_end = comp_end + n;
		const uint8_t* codes = ch->aux0 + (twelve ? sizeof(fls_fsst12_table) : sizeof(fls_fsst_table));
		const fls_fsst12_table* tab12 = reinterpret_cast<const fls_fsst12_table*>(ch->aux0);
		// A warp walks tiles tile, tile + stride, ... of the chunk as a three-stage software pipeline: the entry points of tile
		// k+2 are requested from global memory, the codes of tile k+1 (whose entry points arrived during the previous
		// iteration) are bulk-copied into the other staging buffer, tile k is decoded.  A warp that handled a single tile
		// spent a third of its life waiting for those two round trips.
		const uint32_t n_tiles = (n + tile_strings - 1u) / tile_strings, stride = L.groups_per_rg * kWarps;
		uint32_t tile = g * kWarps + warp;
		if (tile >= n_tiles) {
			return;
		}
		const bool is_rows = L.fuse_rows && FLS_ENC_IS_FSST_ROWS(enc);
		uint4* out_chunk = static_cast<uint4*>(L.out) + (size_t)rg * L.vec_per_rg * 1024u;
		const uint64_t haddr = L.heap_base_addr + hoff;
		uint8_t* cbuf0 = s_dyn + (size_t)warp * (2u * L.cbuf_bytes + L.dbuf_bytes);
		uint8_t* dbuf = cbuf0 + 2u * L.cbuf_bytes;
		struct Tile { // what is LOADED for a tile; everything derived is computed where it is used (so the loads stay in flight)
			uint32_t s0, s1, first, last;   // strings [s0, s1) of the chunk; this lane's strings [first, last]
			uint32_t c0, c1, d0, d1;        // code bytes [c0, c1), decoded bytes [d0, d1) of the tile
			uint32_t cbeg, cend, dbeg;      // where this lane's codes / decoded bytes start, where its codes end
		};
		auto load_tile = [&](uint32_t ti) {
			Tile t;
			t.s0 = ti * tile_strings;
			t.s1 = min(n, t.s0 + tile_strings);
			// tile extents (same address in every lane: one broadcast load each)
			t.c0 = t.s0 ? __ldg(comp_end + t.s0 - 1) : 0u;
			t.d0 = t.s0 ? __ldg(dec_end + t.s0 - 1) : 0u;
			t.c1 = __ldg(comp_end + t.s1 - 1);
			t.d1 = __ldg(dec_end + t.s1 - 1);
			if (L.balance && !twelve) {
				// (experiment, off by default: measured 8-20 % SLOWER than the equal-strings split, the search outweighs the balance)
				// Lanes take equal shares of the tile's CODE BYTES, cut at string boundaries, instead of equal numbers of strings: a
				// lane's work is its code count, and with lane_strings strings apiece the slowest of 32 lanes carried ~25 % more
				// codes than the mean (21.5 active lanes per instruction).  Lane i starts at the first string that begins at or
				// after byte c0 + i * (c1 - c0) / 32 (lower bound over the tile's compressed end offsets: ~10 probes, all but the
				// first few out of L1); its neighbour's start is where it ends.  The heap is one contiguous stream, so any
				// string-aligned cut decodes to the same bytes.
				const uint32_t target = t.c0 + (uint32_t)(((uint64_t)(t.c1 - t.c0) * lane) >> 5);
				uint32_t lo = t.s0, hi = t.s1;
				while (lo < hi) {
					const uint32_t mid = (lo + hi) >> 1;
					const uint32_t start = mid ? __ldg(comp_end + mid - 1) : 0u;
					if (start >= target) {
						hi = mid;
					} else {
						lo = mid + 1u;
					}
				}
				uint32_t nxt = __shfl_down_sync(0xffffffffu, lo, 1);
				if (lane == 31u) {
					nxt = t.s1;
				}
				t.first = lo;
				t.last = nxt ? nxt - 1u : 0u; // unused by the FSST8 decoder
				t.cbeg = lo ? __ldg(comp_end + lo - 1) : 0u;
				t.dbeg = lo ? __ldg(dec_end + lo - 1) : 0u;
				t.cend = nxt ? __ldg(comp_end + nxt - 1) : 0u;
				return t;
			}
			t.first = t.s0 + lane * L.lane_strings;
			t.last = t.first;
			t.cbeg = t.cend = t.dbeg = 0;
			if (t.first < t.s1) {
				t.last = min(t.s1, t.first + L.lane_strings) - 1u;
				t.cbeg = t.first ? __ldg(comp_end + t.first - 1) : 0u;
				t.dbeg = t.first ? __ldg(dec_end + t.first - 1) : 0u;
				t.cend = __ldg(comp_end + t.last);
			}
			return t;
		};
		auto is_staged = [&](const Tile& t) {
			const uint32_t ca = t.c0 & ~15u;
			return ((t.c1 - ca + 15u) & ~15u) <= L.cbuf_bytes && (t.d1 - (t.d0 & ~15u)) + 16u <= L.dbuf_bytes;
		};
		auto stage_codes = [&](const Tile& t, uint32_t b) {
			const uint32_t ca = t.c0 & ~15u, cbytes = (t.c1 - ca + 15u) & ~15u;
			if (is_staged(t) && cbytes && lane == 0) {
				fence_proxy_async(); // the buffer's previous tile was read through the generic proxy
				mbar_expect_tx(&s_bar[2u * warp + b], cbytes);
				tma_load_1d(cbuf0 + b * L.cbuf_bytes, codes + ca, cbytes, &s_bar[2u * warp + b]);
			}
		};
		Tile cur = load_tile(tile);
		stage_codes(cur, 0u);
		Tile nxt = cur;
		if (tile + stride < n_tiles) {
			nxt = load_tile(tile + stride);
		}
		uint32_t phase = 0; // bit b: parity to wait for on staging buffer b
		for (uint32_t k = 0;; ++k) {
			const uint32_t b = k & 1u;
			const bool has_next = tile + stride < n_tiles;
			Tile nn = nxt;
			if (tile + 2u * stride < n_tiles) {
				nn = load_tile(tile + 2u * stride);
			}
			if (has_next) {
				stage_codes(nxt, b ^ 1u);
			}
			const uint32_t s0 = cur.s0, s1 = cur.s1, d0 = cur.d0, d1 = cur.d1, da = cur.d0 & ~15u, ca = cur.c0 & ~15u;
			if (is_staged(cur)) {
				const uint8_t* cbuf = cbuf0 + b * L.cbuf_bytes;
				if (((cur.c1 - ca + 15u) & ~15u) != 0u) {
					mbar_wait(&s_bar[2u * warp + b], (phase >> b) & 1u);
					phase ^= 1u << b;
				}
				if (twelve) {
					if (cur.first < s1) {
						fsst12_decode_lane<false>(cbuf - ca, comp_end, cur.first, cur.last, cur.cbeg, dbuf - da, cur.dbeg, tab12);
					}
				} else {
					fsst_decode_lane<false>(cbuf - ca, cur.cbeg, cur.cend, dbuf - da, cur.dbeg, tab_addr);
				}
				__syncwarp();
				// heap bytes [d0, d1): head bytes up to the first granule boundary, whole granules, tail bytes
				const uint32_t g_lo = (d0 + 15u) & ~15u, g_hi = d1 & ~15u;
				if (g_lo <= g_hi) {
					for (uint32_t q = d0 + lane; q < g_lo; q += 32u) {
						heap[q] = dbuf[q - da];
					}
					for (uint32_t q = g_lo + lane * 16u; q < g_hi; q += 512u) {
						*reinterpret_cast<uint4*>(heap + q) = *reinterpret_cast<const uint4*>(dbuf + (q - da));
					}
					for (uint32_t q = g_hi + lane; q < d1; q += 32u) {
						heap[q] = dbuf[q - da];
					}
				} else { // the whole tile lies inside one granule
					for (uint32_t q = d0 + lane; q < d1; q += 32u) {
						heap[q] = dbuf[q - da];
					}
				}
				if (is_rows) {
					tile_rows<false>(dbuf - da, haddr, dec_end, s0, s1, d0, lane, out_chunk);
				}
				__syncwarp(); // the decode buffer and staging buffer b are free again
			} else {
				// oversize tile (very long strings): decode straight to the heap
				if (twelve) {
					if (cur.first < s1) {
						fsst12_decode_lane<true>(codes, comp_end, cur.first, cur.last, cur.cbeg, heap, cur.dbeg, tab12);
					}
				} else {
					fsst_decode_lane<true>(codes, cur.cbeg, cur.cend, heap, cur.dbeg, tab_addr);
				}
				if (is_rows) {
					__threadfence_block();
					__syncwarp();
					tile_rows<true>(heap, haddr, dec_end, s0, s1, d0, lane, out_chunk);
				}
			}
			if (!has_next) {
				break;
			}
			cur = nxt;
			nxt = nn;
			tile += stride;
		}
		return;
	}
	// plain bytes: STR_UNCOMPRESSED / DICT_STR / STR_CONSTANT
	const uint64_t bytes = ch->const_val;
	const uint64_t n_gran = (bytes + 15u) >> 4; // the heap slice is padded to 16 bytes, the source segment has 128 bytes of slack
	const uint64_t per = (n_gran + L.groups_per_rg - 1) / L.groups_per_rg;
	const uint64_t g0 = (uint64_t)g * per, g1 = min(n_gran, g0 + per);
	const uint4* src = reinterpret_cast<const uint4*>(ch->aux0);
	uint4* dst = reinterpret_cast<uint4*>(heap);
	for (uint64_t i = g0 + threadIdx.x; i < g1; i += blockDim.x) {
		dst[i] = __ldg(src + i);
	}
	if (L.fuse_keys && enc == FLS_ENC_DICT_STR) { // the keys' string_t: built from the source bytes the slice is a copy of
		const uint32_t n = ch->aux_count;
		const uint32_t* ends = reinterpret_cast<const uint32_t*>(ch->aux1);
		uint4* out = static_cast<uint4*>(L.out) + (L.key_prefix[(size_t)rg * L.chunk_stride] - L.key_prefix0);
		for (uint32_t k = g * blockDim.x + threadIdx.x; k < n; k += L.groups_per_rg * blockDim.x) {
			const uint32_t start = k ? __ldg(ends + k - 1) : 0u, end = __ldg(ends + k);
			out[k] = make_string_t<false>(ch->aux0, L.heap_base_addr + hoff, start, end - start);
		}
	}
}

template <int IT>
__device__ __forceinline__ void codes_to_tile(const FlsChunkDev* ch, uint32_t vi, uint32_t lane, uint32_t* tile) {
	const uint4* mp = reinterpret_cast<const uint4*>(ch->meta + vi);
	const uint4 m0 = __ldg(mp), m1 = __ldg(mp + 1);
	const uint32_t base = m0.x, W = (m1.x >> 16) & 0xffu;
	const uint32_t* blob = reinterpret_cast<const uint32_t*>(ch->data + (size_t)m0.z * 16);
	uint32_t v[32];
	unpack_small_dispatch<IT>((int)W, blob, (int)lane, v);
	const uint32_t n_keys = ch->aux_count;
#pragma unroll
	for (int i = 0; i < 32; ++i) {
		tile[value_pos<IT>((int)lane, i)] = min(v[i] + base, n_keys - 1u); // clamp: never index outside the dictionary
	}
}

__device__ __forceinline__ void vector_codes(const FlsChunkDev* ch, uint32_t vi, uint32_t lane, uint32_t* tile) {
	const uint32_t ib = ch->idx_bits;
	if (ib == 8) {
		codes_to_tile<8>(ch, vi, lane, tile);
	} else if (ib == 16) {
		codes_to_tile<16>(ch, vi, lane, tile);
	} else {
		codes_to_tile<32>(ch, vi, lane, tile);
	}
	__syncwarp();
}

// VALUE encodings (fls_format.h FLS_ENC_IS_STR_VALUES): tile[row] = which of the chunk's value strings row `row` of vector vi is.
//   RLE_STR (rle_map_kernel.hpp:26-61): the DELTA-coded u16 run index, unpacked and summed per FastLanes lane exactly like the
//     numeric RLE index (process_rle): a thread owns lanes 2*lane, 2*lane + 1, whose 16 rows are 16 consecutive logical positions
//     starting at chain_pos; value = first run of the vector + index
//   CROSS_RLE_STR (cross_rle_kernel.hpp:55-97): binary search of the row's position in the run starts (prefix sums built at
//     stage time) instead of the reference's rescan of the run lengths per vector
//   FREQUENCY_STR (frequency_kernel.hpp:23-69): every row is the vector's frequent string, then the exceptions are patched in
__device__ __forceinline__ int chain_pos(int fl_lane) { // logical start of the 16-row chain of 16-bit lane fl_lane
	return (fl_lane & 15) * 64 + fl_order((fl_lane >> 4) & 7) * 8;
}

__device__ __forceinline__ void vector_value_index(const FlsChunkDev* ch, uint32_t enc, uint32_t vi, uint32_t lane, uint32_t* tile) {
	const uint4* mp = reinterpret_cast<const uint4*>(ch->meta + vi);
	const uint4 m0 = __ldg(mp), m1 = __ldg(mp + 1); // {base lo, base hi, data_off16, aux_off}, {aux_cnt | width << 16 ..}
	const uint32_t first = m0.w, cnt = m1.x & 0xffffu, n_vals = ch->aux_count;
	if (enc == FLS_ENC_RLE_STR) {
		const uint32_t W = (m1.x >> 16) & 0xffu, dmin = m0.x;
		const uint32_t* blob = reinterpret_cast<const uint32_t*>(ch->data + (size_t)m0.z * 16);
		uint32_t v[32];
		unpack_small_dispatch<16>((int)W, blob + 32, (int)lane, v);
		const uint32_t bw = __ldg(blob + lane);
#pragma unroll
		for (int h = 0; h < 2; ++h) {
			uint32_t acc = (bw >> (16 * h)) & 0xffffu;
			const int p0 = chain_pos(2 * (int)lane + h);
#pragma unroll
			for (int row = 0; row < 16; ++row) {
				acc += v[2 * row + h] + dmin;
				tile[p0 + row] = min(first + min(acc & 0xffffu, cnt - 1u), n_vals - 1u); // clamped: never outside the value strings
			}
		}
	} else if (enc == FLS_ENC_CROSS_RLE_STR) {
		const uint32_t* starts = reinterpret_cast<const uint32_t*>(ch->data);
		const uint32_t g0 = vi * 1024u;
		uint32_t lo_r = 0, hi_r = n_vals; // first run of this vector: largest r with starts[r] <= g0
		while (hi_r - lo_r > 1) {
			const uint32_t mid = (lo_r + hi_r) >> 1;
			if (__ldg(starts + mid) <= g0) {
				lo_r = mid;
			} else {
				hi_r = mid;
			}
		}
#pragma unroll 1
		for (uint32_t it = 0; it < 32u; ++it) {
			const uint32_t gi = g0 + it * 32u + lane;
			uint32_t a = lo_r, b = min(n_vals, lo_r + 1025u);
			while (b - a > 1) {
				const uint32_t mid = (a + b) >> 1;
				if (__ldg(starts + mid) <= gi) {
					a = mid;
				} else {
					b = mid;
				}
			}
			tile[it * 32u + lane] = a;
		}
	} else { // FREQUENCY_STR
		const uint32_t freq = min(first, n_vals - 1u);
#pragma unroll 4
		for (uint32_t it = 0; it < 32u; ++it) {
			tile[it * 32u + lane] = freq;
		}
		__syncwarp();
		const uint16_t* pos = reinterpret_cast<const uint16_t*>(ch->data + (size_t)m0.z * 16);
		for (uint32_t k = lane; k < cnt; k += 32u) {
			tile[__ldg(pos + k) & 1023u] = min(first + 1u + k, n_vals - 1u);
		}
	}
	__syncwarp();
}

// one warp per vector: 1024 string_t (flat form), or the u32 codes (selection vector) when L.codes_only
__device__ __forceinline__ void str_rows_body(const FlsStrLaunch& L) {
	__shared__ uint32_t s_codes[kWarps][1024];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const uint32_t v = blockIdx.x * kWarps + warp;
	if (v >= L.n_rg * L.vec_per_rg) {
		return;
	}
	const uint32_t rg = v / L.vec_per_rg, vi = v - rg * L.vec_per_rg;
	const FlsChunkDev* ch = L.chunks + (size_t)rg * L.chunk_stride;
	if (vi >= ch->n_vec) {
		return;
	}
	const uint32_t enc = ch->encoding;
	const bool dict = FLS_ENC_IS_STR_DICT(enc);
	if (L.fuse_rows && FLS_ENC_IS_FSST_ROWS(enc)) {
		return; // the heap kernel has built these rows from its staging buffer
	}
	if (L.key_flags) { // filter mode: code -> flag of its key -> selection mask (host: every chunk of the range is a dictionary)
		if (!dict) {
			return;
		}
		const uint32_t nt = __ldg(L.n_tuples + rg), first = vi * 1024u;
		const bool keep = L.rg_keep == nullptr || __ldg(L.rg_keep + rg) != 0;
		const uint32_t valid = !keep || nt <= first ? 0u : (nt - first < 1024u ? nt - first : 1024u);
		uint32_t* tile = s_codes[warp];
		vector_codes(ch, vi, lane, tile);
		const uint8_t* fl = L.key_flags + (L.key_prefix[(size_t)rg * L.chunk_stride] - L.key_prefix0);
		uint32_t mine = 0;
#pragma unroll 4
		for (uint32_t it = 0; it < 32u; ++it) {
			const uint32_t row = it * 32u + lane;
			const bool pass = row < valid && __ldg(fl + tile[row]) != 0;
			const uint32_t word = __ballot_sync(0xffffffffu, pass);
			if (lane == it) {
				mine = word;
			}
		}
		uint32_t* mw = L.filter_mask + (size_t)v * 32u + lane;
		*mw = L.filter_and ? (*mw & mine) : mine;
		return;
	}
	if (L.codes_only) {
		if (dict) {
			uint32_t* tile = s_codes[warp];
			vector_codes(ch, vi, lane, tile);
			uint4* out = reinterpret_cast<uint4*>(static_cast<uint32_t*>(L.out) + (size_t)v * 1024u);
			const uint4* t4 = reinterpret_cast<const uint4*>(tile);
#pragma unroll
			for (int it = 0; it < 8; ++it) {
				out[it * 32 + lane] = t4[it * 32 + lane];
			}
		}
		return;
	}
	const uint64_t hoff = L.heap_prefix[(size_t)rg * L.chunk_stride] - L.heap_prefix0;
	const uint8_t* hp = L.heap + hoff;
	const uint64_t haddr = L.heap_base_addr + hoff;
	uint4* out = static_cast<uint4*>(L.out) + (size_t)v * 1024u;
	const uint32_t* ends = reinterpret_cast<const uint32_t*>(ch->aux1) + (is_fsst(enc) ? ch->aux_count : 0u); // decoded end offsets
	if (enc == FLS_ENC_STR_CONSTANT) {
		const uint4 s = make_string_t<false>(hp, haddr, 0, ch->aux_count);
#pragma unroll 4
		for (int it = 0; it < 32; ++it) {
			out[it * 32 + lane] = s;
		}
	} else if (dict || FLS_ENC_IS_STR_VALUES(enc)) {
		uint32_t* tile = s_codes[warp];
		if (dict) {
			vector_codes(ch, vi, lane, tile);
		} else {
			vector_value_index(ch, enc, vi, lane, tile);
		}
#pragma unroll 2
		for (int it = 0; it < 32; ++it) {
			const uint32_t k = tile[it * 32 + lane];
			const uint32_t start = k ? __ldg(ends + k - 1) : 0u, end = __ldg(ends + k);
			out[it * 32 + lane] = make_string_t<false>(hp, haddr, start, end - start);
		}
	} else { // STR_UNCOMPRESSED, FSST: one string per row
		const uint32_t i0 = vi * 1024u;
		uint32_t carry = i0 ? __ldg(ends + i0 - 1) : 0u; // end of the string before this warp step
#pragma unroll 2
		for (int it = 0; it < 32; ++it) {
			const uint32_t end = __ldg(ends + i0 + it * 32 + lane);
			uint32_t start = __shfl_up_sync(0xffffffffu, end, 1);
			if (lane == 0) {
				start = carry;
			}
			carry = __shfl_sync(0xffffffffu, end, 31);
			out[it * 32 + lane] = make_string_t<false>(hp, haddr, start, end - start);
		}
	}
}

// dictionary form: string_t of every key of the dictionary chunks of the range (dictionary_kernel.hpp:60-78,
// fsst_dict_kernel.hpp:54-80); chunk rg's keys start at key_prefix[rg] - key_prefix0.  grid: n_rg * groups_per_rg CTAs.
__device__ __forceinline__ void str_keys_body(const FlsStrLaunch& L) {
	const uint32_t rg = blockIdx.x / L.groups_per_rg, g = blockIdx.x - rg * L.groups_per_rg;
	const FlsChunkDev* ch = L.chunks + (size_t)rg * L.chunk_stride;
	const uint32_t enc = ch->encoding;
	if (!FLS_ENC_IS_STR_DICT(enc)) {
		return;
	}
	const uint32_t n = ch->aux_count;
	const uint32_t* ends = reinterpret_cast<const uint32_t*>(ch->aux1) + (is_fsst(enc) ? n : 0u);
	const uint64_t hoff = L.heap_prefix[(size_t)rg * L.chunk_stride] - L.heap_prefix0;
	uint4* out = static_cast<uint4*>(L.out) + (L.key_prefix[(size_t)rg * L.chunk_stride] - L.key_prefix0);
	for (uint32_t k = g * blockDim.x + threadIdx.x; k < n; k += L.groups_per_rg * blockDim.x) {
		const uint32_t start = k ? __ldg(ends + k - 1) : 0u, end = __ldg(ends + k);
		out[k] = make_string_t<false>(L.heap + hoff, L.heap_base_addr + hoff, start, end - start);
	}
}

// The kernels proper.  With L.pdl (launch chaining, see launch_family in flsgpu_device.cu) a grid releases its successor at
// once and waits for its predecessor before it exits: the heap kernel reads nothing but staged bytes, so it may overlap the
// decode launch in front of it; the codes (selection vector) launch is independent of the heap / keys launches of the same
// dictionary decode and overlaps them; the keys launch needs the heap complete and is an ordinary launch.
__global__ void __launch_bounds__(kWarps * 32) fls_str_heap_kernel(const FlsStrLaunch L) {
	if (L.pdl) {
		griddep_launch_dependents();
	}
	str_heap_body(L);
	if (L.pdl) {
		griddep_wait();
	}
}

__global__ void __launch_bounds__(kWarps * 32) fls_str_rows_kernel(const FlsStrLaunch L) {
	if (L.pdl) {
		griddep_launch_dependents();
	}
	str_rows_body(L);
	if (L.pdl) {
		griddep_wait();
	}
}

__global__ void __launch_bounds__(256) fls_str_keys_kernel(const FlsStrLaunch L) {
	if (L.pdl) {
		griddep_launch_dependents();
	}
	str_keys_body(L);
}

// FilterExecutor::Apply on a string column (reference filter_executor.cpp:19-60 -> ColumnSegment::FilterSelection with a
// ConstantFilter): compare every decoded string_t of the slab with the constant the way DuckDB does (string_type.hpp:155-217:
// memcmp over the common length, the shorter string first on a tie) and emit the selection as mask words.  One warp per
// vector: 32 coalesced 512-byte row loads; most rows are decided by the 4-byte prefix (zero padded in string_t, so a
// big-endian integer comparison of two different prefixes is the bytewise order), the rest walk their bytes (inlined:
// the row itself; longer: the heap the rows point into).
__device__ __forceinline__ int compare_string_row(const uint4& s, const uint8_t* row_bytes, const FlsStrFilterLaunch& L) {
	const uint32_t pa = __byte_perm(s.y, 0u, 0x0123), pb = __byte_perm(L.const_prefix, 0u, 0x0123);
	if (pa != pb) {
		return pa < pb ? -1 : 1;
	}
	const uint32_t la = s.x, lb = L.const_len, m = la < lb ? la : lb;
	if (m > 4u) {
		const uint8_t* p = la <= 12u ? row_bytes + 4 : reinterpret_cast<const uint8_t*>(((unsigned long long)s.w << 32) | s.z);
		for (uint32_t k = 4; k < m; ++k) {
			const uint32_t a = p[k], b = __ldg(L.const_data + k);
			if (a != b) {
				return a < b ? -1 : 1;
			}
		}
	}
	return la < lb ? -1 : la > lb ? 1 : 0;
}

__device__ __forceinline__ bool passes_cmp(uint32_t op, int c) { // flsgpu_cmp: EQ NE LT LE GT GE
	switch (op) {
	case 0: return c == 0;
	case 1: return c != 0;
	case 2: return c < 0;
	case 3: return c <= 0;
	case 4: return c > 0;
	default: return c >= 0;
	}
}

__global__ void __launch_bounds__(kWarps * 32) fls_str_filter_kernel(const FlsStrFilterLaunch L) {
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const uint32_t v = blockIdx.x * kWarps + warp;
	if (v >= L.n_rg * L.vec_per_rg) {
		return;
	}
	const uint32_t rg = v / L.vec_per_rg, vi = v - rg * L.vec_per_rg;
	const uint32_t nt = __ldg(L.n_tuples + rg);
	const bool keep = L.rg_keep == nullptr || __ldg(L.rg_keep + rg) != 0;
	const uint32_t first = vi * 1024u;
	const uint32_t valid = !keep || nt <= first ? 0u : (nt - first < 1024u ? nt - first : 1024u);
	const uint4* rows = static_cast<const uint4*>(L.rows) + (size_t)v * 1024u;
	uint32_t mine = 0;
	for (uint32_t it = 0; it * 32u < valid; ++it) {
		const uint32_t row = it * 32u + lane;
		bool pass = false;
		if (row < valid) {
			const uint4 s = __ldg(rows + row);
			pass = passes_cmp(L.op, compare_string_row(s, reinterpret_cast<const uint8_t*>(rows + row), L));
		}
		const uint32_t word = __ballot_sync(0xffffffffu, pass);
		if (lane == it) {
			mine = word;
		}
	}
	uint32_t* mw = L.mask + (size_t)v * 32u + lane;
	*mw = L.and_into ? (*mw & mine) : mine;
}

// dictionary form: the comparison once per key of the range's dictionaries (a few thousand keys instead of every row)
__global__ void __launch_bounds__(256) fls_str_key_flags_kernel(const FlsStrFilterLaunch L) {
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= L.n_keys) {
		return;
	}
	const uint4* rows = static_cast<const uint4*>(L.rows);
	const uint4 s = __ldg(rows + k);
	const int c = compare_string_row(s, reinterpret_cast<const uint8_t*>(rows + k), L);
	L.key_flags_out[k] = passes_cmp(L.op, c) ? 1 : 0;
}

// <<<>>> with programmatic stream serialization when `chained` (the grid may start while its predecessor still runs)
cudaError_t launch_chained(void (*kernel)(const FlsStrLaunch), const FlsStrLaunch& L, uint32_t grid, uint32_t block, uint32_t smem,
                           cudaStream_t stream, bool chained) {
	cudaLaunchConfig_t cfg {};
	cfg.gridDim = dim3(grid);
	cfg.blockDim = dim3(block);
	cfg.dynamicSmemBytes = smem;
	cfg.stream = stream;
	cudaLaunchAttribute attr {};
	attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr.val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = &attr;
	cfg.numAttrs = chained ? 1 : 0;
	return cudaLaunchKernelEx(&cfg, kernel, L);
}

} // namespace

cudaError_t fls_str_launch_key_flags(FlsStrFilterLaunch L, cudaStream_t stream) {
	if (L.n_keys == 0) {
		return cudaSuccess;
	}
	fls_str_key_flags_kernel<<<(L.n_keys + 255u) / 256u, 256, 0, stream>>>(L);
	return cudaGetLastError();
}

cudaError_t fls_str_launch_filter(FlsStrFilterLaunch L, cudaStream_t stream) {
	const uint32_t n_vec = L.n_rg * L.vec_per_rg;
	fls_str_filter_kernel<<<(n_vec + kWarps - 1) / kWarps, kWarps * 32, 0, stream>>>(L);
	return cudaGetLastError();
}

cudaError_t fls_str_launch_heap(FlsStrLaunch L, cudaStream_t stream) {
	// 8 warp tiles of 32 * lane_strings strings per CTA; groups_per_rg (CTAs per chunk) comes from the launcher, which knows
	// the largest chunk of the range
	if (L.lane_strings == 0) {
		L.lane_strings = 8;
	}
	if (L.groups_per_rg == 0) {
		const uint32_t per_cta = kWarps * 32u * L.lane_strings;
		L.groups_per_rg = (L.vec_per_rg * 1024u + per_cta - 1u) / per_cta;
	}
	const uint32_t smem = kWarps * (2u * L.cbuf_bytes + L.dbuf_bytes); // two staging buffers for codes + one decode buffer per warp
	static bool attr_set[64] = {}; // per device; benign race: the attribute is idempotent
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 0 || dev >= 64 || !attr_set[dev]) {
		const cudaError_t e = cudaFuncSetAttribute(fls_str_heap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
		if (e != cudaSuccess) {
			return e;
		}
		if (dev >= 0 && dev < 64) {
			attr_set[dev] = true;
		}
	}
	return launch_chained(fls_str_heap_kernel, L, L.n_rg * L.groups_per_rg, kWarps * 32, smem, stream, L.pdl != 0);
}

cudaError_t fls_str_launch_rows(FlsStrLaunch L, cudaStream_t stream) {
	const uint32_t n_vec = L.n_rg * L.vec_per_rg;
	// only the selection-vector form is independent of the launches in front of it (flat rows read the heap, filter mode the flags)
	return launch_chained(fls_str_rows_kernel, L, (n_vec + kWarps - 1) / kWarps, kWarps * 32, 0, stream,
	                      L.pdl != 0 && L.codes_only != 0 && L.key_flags == nullptr);
}

cudaError_t fls_str_launch_keys(FlsStrLaunch L, uint32_t max_keys_per_chunk, cudaStream_t stream) {
	L.groups_per_rg = (uint32_t)((max_keys_per_chunk + 255u) / 256u);
	L.groups_per_rg = L.groups_per_rg < 1u ? 1u : L.groups_per_rg > 64u ? 64u : L.groups_per_rg;
	fls_str_keys_kernel<<<L.n_rg * L.groups_per_rg, 256, 0, stream>>>(L);
	return cudaGetLastError();
}
