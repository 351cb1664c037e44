// fls_fsst.cpp — FSST symbol-table trainer + greedy encoder (see fls_fsst.hpp).
#include "fls_fsst.hpp"

#include <algorithm>
#include <cstring>

namespace flshost {

namespace {

inline uint64_t load_prefix(const uint8_t* s, uint32_t len) { // first `len` (<= 8) bytes, little endian, zero padded
	uint64_t v = 0;
	std::memcpy(&v, s, len);
	return v;
}

constexpr uint64_t kLenMask[9] = {0ull,
                                  0xffull,
                                  0xffffull,
                                  0xffffffull,
                                  0xffffffffull,
                                  0xffffffffffull,
                                  0xffffffffffffull,
                                  0xffffffffffffffull,
                                  ~0ull};

struct Cand {
	uint64_t value;
	uint32_t len;
	uint64_t gain;
};

// (bytes, length) -> occurrences, for the trainers: open addressing, grows at half load.  The order candidates come out in
// does not matter: the selection below sorts them by a total order.
struct SymbolCounts {
	std::vector<uint64_t> value, count;
	std::vector<uint8_t> len; // 0 = empty slot
	uint64_t used = 0;

	explicit SymbolCounts(uint32_t log2_slots) : value(1ull << log2_slots), count(1ull << log2_slots), len(1ull << log2_slots) {
	}
	void clear() {
		std::fill(len.begin(), len.end(), (uint8_t)0);
		used = 0;
	}
	static uint64_t hash(uint64_t v, uint32_t l) {
		return ((v ^ ((uint64_t)l << 59)) * 0x9E3779B97F4A7C15ull) >> 24;
	}
	void add(uint64_t v, uint32_t l) {
		const uint64_t mask = len.size() - 1;
		uint64_t h = hash(v, l) & mask;
		while (len[h]) {
			if (len[h] == l && value[h] == v) {
				++count[h];
				return;
			}
			h = (h + 1) & mask;
		}
		value[h] = v;
		len[h] = (uint8_t)l;
		count[h] = 1;
		if (++used * 2 > len.size()) {
			grow();
		}
	}
	void grow() {
		SymbolCounts bigger(0);
		bigger.value.assign(len.size() * 2, 0);
		bigger.count.assign(len.size() * 2, 0);
		bigger.len.assign(len.size() * 2, 0);
		const uint64_t mask = bigger.len.size() - 1;
		for (size_t i = 0; i < len.size(); ++i) {
			if (len[i]) {
				uint64_t h = hash(value[i], len[i]) & mask;
				while (bigger.len[h]) {
					h = (h + 1) & mask;
				}
				bigger.value[h] = value[i];
				bigger.len[h] = len[i];
				bigger.count[h] = count[i];
			}
		}
		bigger.used = used;
		*this = std::move(bigger);
	}
};

} // namespace

// Symbols of two or more bytes, grouped by their first two bytes and ordered longest first inside a group: the longest
// symbol that matches at a position is the first hit of one short list (a handful of entries on text), where probing a hash
// per candidate length cost up to seven lookups per position.
void FsstPrefixIndex::build(const uint64_t* symbol, const uint8_t* len, uint32_t first_code, uint32_t n_codes) {
	bucket.assign(1u << 16, 0u);
	order.clear();
	max_len = 0;
	std::vector<uint32_t> fill(1u << 16, 0u);
	uint32_t n_mid = 0, n_full = 0;
	for (uint32_t c = first_code; c < first_code + n_codes; ++c) {
		max_len = std::max<uint8_t>(max_len, len[c]);
		if (len[c] >= 2 && len[c] < 8) {
			++fill[symbol[c] & 0xffffu];
			++n_mid;
		}
		n_full += len[c] == 8;
	}
	// 8-byte symbols (half of a table trained on text): one probe with the whole word
	uint32_t log2 = 4;
	while ((1u << log2) < 4 * n_full) {
		++log2;
	}
	full.assign(1u << log2, 0);
	full_shift = 64 - log2;
	for (uint32_t c = first_code; c < first_code + n_codes; ++c) {
		if (len[c] == 8) {
			uint32_t h = (uint32_t)((symbol[c] * 0x9E3779B97F4A7C15ull) >> full_shift);
			while (full[h]) {
				h = (h + 1) & ((uint32_t)full.size() - 1);
			}
			full[h] = (uint16_t)(c + 1);
		}
	}
	order.resize(n_mid);
	uint32_t start = 0;
	for (uint32_t k = 0; k < (1u << 16); ++k) {
		if (fill[k]) {
			bucket[k] = (start << 16) | fill[k];
			start += fill[k];
			fill[k] = 0;
		}
	}
	for (uint32_t c = first_code; c < first_code + n_codes; ++c) {
		if (len[c] >= 2 && len[c] < 8) {
			const uint32_t k = (uint32_t)(symbol[c] & 0xffffu);
			order[(bucket[k] >> 16) + fill[k]++] = (uint16_t)c;
		}
	}
	for (uint32_t k = 0; k < (1u << 16); ++k) {
		if ((bucket[k] & 0xffffu) > 1) {
			uint16_t* o = order.data() + (bucket[k] >> 16);
			std::sort(o, o + (bucket[k] & 0xffffu), [&](uint16_t a, uint16_t b) { return len[a] != len[b] ? len[a] > len[b] : a < b; });
		}
	}
}

int FsstPrefixIndex::find(const uint64_t* symbol, const uint8_t* len, const uint8_t* s, uint64_t remaining, uint32_t* len_out) const {
	if (remaining < 2) {
		return -1;
	}
	uint64_t w = 0;
	if (remaining >= 8) {
		std::memcpy(&w, s, 8);
		for (uint32_t h = (uint32_t)((w * 0x9E3779B97F4A7C15ull) >> full_shift); full[h]; h = (h + 1) & ((uint32_t)full.size() - 1)) {
			if (symbol[full[h] - 1u] == w) {
				*len_out = 8;
				return (int)full[h] - 1;
			}
		}
	} else {
		std::memcpy(&w, s, remaining);
	}
	const uint32_t e = bucket[(uint32_t)(w & 0xffffu)];
	if (!e) {
		return -1;
	}
	const uint16_t* o = order.data() + (e >> 16);
	for (uint32_t k = 0, n = e & 0xffffu; k < n; ++k) {
		const uint32_t c = o[k], l = len[c];
		if (l <= remaining && (w & kLenMask[l]) == symbol[c]) {
			*len_out = l;
			return (int)c;
		}
	}
	return -1;
}

void FsstEncoder::rebuild_lookup() {
	std::memset(byte_code, 0, sizeof(byte_code));
	for (uint32_t c = 0; c < n_symbols; ++c) {
		if (table.len[c] == 1) {
			byte_code[table.symbol[c] & 0xff] = (uint16_t)(c + 1);
		}
	}
	index.build(table.symbol, table.len, 0, n_symbols);
}

// longest symbol matching at s; returns its code or -1
int FsstEncoder::find(const uint8_t* s, uint64_t remaining, uint32_t* len_out) const {
	const int c = index.find(table.symbol, table.len, s, remaining, len_out);
	if (c >= 0) {
		return c;
	}
	if (remaining && byte_code[s[0]]) {
		*len_out = 1;
		return byte_code[s[0]] - 1;
	}
	return -1;
}

void FsstEncoder::train(const uint64_t* offsets, const uint8_t* bytes, uint64_t n) {
	// sample ~64 KiB of whole strings, spread over the input
	std::vector<std::pair<const uint8_t*, uint64_t>> sample;
	const uint64_t total = n ? offsets[n] - offsets[0] : 0;
	const uint64_t budget = 64u << 10;
	const uint64_t stride = total > budget ? std::max<uint64_t>(1, n / std::max<uint64_t>(1, n * budget / total)) : 1;
	uint64_t taken = 0;
	for (uint64_t i = 0; i < n && taken < 2 * budget; i += stride) {
		const uint64_t len = offsets[i + 1] - offsets[i];
		if (len) {
			sample.emplace_back(bytes + offsets[i], len);
			taken += len;
		}
	}
	n_symbols = 0;
	std::memset(&table, 0, sizeof(table));
	rebuild_lookup();
	SymbolCounts count(15);
	for (int round = 0; round < 5; ++round) {
		count.clear();
		for (const auto& sv : sample) {
			const uint8_t* s = sv.first;
			const uint64_t len = sv.second;
			uint64_t prev_v = 0;
			uint32_t prev_len = 0;
			for (uint64_t p = 0; p < len;) {
				uint32_t l = 1;
				const int c = find(s + p, len - p, &l);
				if (c < 0) {
					l = 1;
				}
				const uint64_t v = load_prefix(s + p, l);
				count.add(v, l);
				if (prev_len && prev_len + l <= 8) { // candidate: the two adjacent symbols glued together
					count.add(prev_v | (v << (8 * prev_len)), prev_len + l);
				}
				prev_v = v;
				prev_len = l;
				p += l;
			}
		}
		std::vector<Cand> cands;
		cands.reserve(count.used);
		for (size_t i = 0; i < count.len.size(); ++i) {
			if (!count.len[i]) {
				continue;
			}
			// a single byte that is not a symbol costs 2 output bytes (escape), so a 1-byte symbol saves 1 per use like a
			// 2-byte symbol does; longer symbols save len - 1
			const uint64_t n = count.count[i], l = count.len[i];
			const uint64_t gain = n * (l == 1 ? 1 : l - 1) + (l == 1 ? n / 2 : 0);
			cands.push_back({count.value[i], (uint32_t)l, gain});
		}
		const size_t keep = std::min<size_t>(255, cands.size());
		std::partial_sort(cands.begin(), cands.begin() + keep, cands.end(), [](const Cand& a, const Cand& b) {
			return a.gain != b.gain ? a.gain > b.gain : a.len != b.len ? a.len > b.len : a.value < b.value;
		});
		n_symbols = (uint32_t)keep;
		std::memset(&table, 0, sizeof(table));
		for (size_t i = 0; i < keep; ++i) {
			table.symbol[i] = cands[i].value;
			table.len[i] = (uint8_t)cands[i].len;
		}
		rebuild_lookup();
	}
	// unused codes decode to one zero byte: harmless and keeps every len in 1..8 (the escape slot included)
	for (uint32_t c = n_symbols; c < 256; ++c) {
		table.len[c] = 1;
	}
}

void FsstEncoder::compress(const uint8_t* s, uint64_t len, std::vector<uint8_t>& out) const {
	for (uint64_t p = 0; p < len;) {
		uint32_t l = 1;
		const int c = find(s + p, len - p, &l);
		if (c < 0) {
			out.push_back((uint8_t)FLS_FSST_ESC);
			out.push_back(s[p]);
			++p;
		} else {
			out.push_back((uint8_t)c);
			p += l;
		}
	}
}

uint64_t fsst_decoded_length(const fls_fsst_table& table, const uint8_t* codes, uint64_t n) {
	uint64_t out = 0;
	for (uint64_t p = 0; p < n; ++p) {
		if (codes[p] == FLS_FSST_ESC) {
			if (++p >= n) {
				return ~0ull; // escape without its literal
			}
			++out;
		} else {
			out += table.len[codes[p]];
		}
	}
	return out;
}

// ---- FSST12 -----------------------------------------------------------------------------------------------------------
void Fsst12Encoder::rebuild_lookup() {
	index.build(table.symbol, table.len, 256, n_symbols);
}

int Fsst12Encoder::find(const uint8_t* s, uint64_t remaining, uint32_t* len_out) const {
	return index.find(table.symbol, table.len, s, remaining, len_out);
}

void Fsst12Encoder::train(const uint64_t* offsets, const uint8_t* bytes, uint64_t n) {
	std::vector<std::pair<const uint8_t*, uint64_t>> sample;
	const uint64_t total = n ? offsets[n] - offsets[0] : 0;
	const uint64_t budget = 256u << 10; // 3840 symbols want a larger sample than FSST's 255
	const uint64_t stride = total > budget ? std::max<uint64_t>(1, n / std::max<uint64_t>(1, n * budget / total)) : 1;
	uint64_t taken = 0;
	for (uint64_t i = 0; i < n && taken < 2 * budget; i += stride) {
		const uint64_t len = offsets[i + 1] - offsets[i];
		if (len) {
			sample.emplace_back(bytes + offsets[i], len);
			taken += len;
		}
	}
	n_symbols = 0;
	std::memset(&table, 0, sizeof(table));
	for (uint32_t c = 0; c < 256; ++c) { // codes < 256: the byte itself
		table.symbol[c] = c;
		table.len[c] = 1;
	}
	rebuild_lookup();
	SymbolCounts count(17);
	for (int round = 0; round < 5; ++round) {
		count.clear();
		for (const auto& sv : sample) {
			const uint8_t* s = sv.first;
			const uint64_t len = sv.second;
			uint64_t prev_v = 0;
			uint32_t prev_len = 0;
			for (uint64_t p = 0; p < len;) {
				uint32_t l = 1;
				if (find(s + p, len - p, &l) < 0) {
					l = 1;
				}
				const uint64_t v = load_prefix(s + p, l);
				if (l >= 2) {
					count.add(v, l);
				}
				if (prev_len && prev_len + l <= 8) {
					count.add(prev_v | (v << (8 * prev_len)), prev_len + l);
				}
				prev_v = v;
				prev_len = l;
				p += l;
			}
		}
		std::vector<Cand> cands;
		cands.reserve(count.used);
		for (size_t i = 0; i < count.len.size(); ++i) {
			if (count.len[i] && count.count[i] >= 2) { // a symbol seen once saves nothing worth a table slot
				cands.push_back({count.value[i], (uint32_t)count.len[i], count.count[i] * (count.len[i] - 1u)});
			}
		}
		const size_t keep = std::min<size_t>(4096 - 256, cands.size());
		std::partial_sort(cands.begin(), cands.begin() + keep, cands.end(), [](const Cand& a, const Cand& b) {
			return a.gain != b.gain ? a.gain > b.gain : a.len != b.len ? a.len > b.len : a.value < b.value;
		});
		n_symbols = (uint32_t)keep;
		for (uint32_t c = 256; c < 4096; ++c) {
			table.symbol[c] = 0;
			table.len[c] = 0;
		}
		for (size_t i = 0; i < keep; ++i) {
			table.symbol[256 + i] = cands[i].value;
			table.len[256 + i] = (uint8_t)cands[i].len;
		}
		rebuild_lookup();
	}
	for (uint32_t c = 256 + n_symbols; c < 4096; ++c) { // unused codes: one zero byte (keeps every len in 1..8)
		table.len[c] = 1;
	}
}

void Fsst12Encoder::compress(const uint8_t* s, uint64_t len, std::vector<uint8_t>& out) const {
	uint32_t pending = 0;
	bool have = false;
	for (uint64_t p = 0; p < len;) {
		uint32_t l = 1;
		int c = find(s + p, len - p, &l);
		if (c < 0) {
			c = s[p];
			l = 1;
		}
		p += l;
		if (!have) {
			pending = (uint32_t)c;
			have = true;
		} else { // two codes -> 3 bytes
			out.push_back((uint8_t)(pending & 0xff));
			out.push_back((uint8_t)((pending >> 8) | (((uint32_t)c & 15u) << 4)));
			out.push_back((uint8_t)((uint32_t)c >> 4));
			have = false;
		}
	}
	if (have) { // odd code count: the last code takes 2 bytes
		out.push_back((uint8_t)(pending & 0xff));
		out.push_back((uint8_t)(pending >> 8));
	}
}

uint64_t fsst12_decoded_length(const fls_fsst12_table& table, const uint8_t* codes, uint64_t n) {
	uint64_t out = 0, p = 0;
	for (; p + 3 <= n; p += 3) {
		out += table.len[codes[p] | ((codes[p + 1] & 15u) << 8)];
		out += table.len[(codes[p + 1] >> 4) | ((uint32_t)codes[p + 2] << 4)];
	}
	if (n - p == 2) {
		if (codes[p + 1] >> 4) {
			return ~0ull; // padding bits of the final code must be zero
		}
		out += table.len[codes[p] | ((codes[p + 1] & 15u) << 8)];
	} else if (n - p == 1) {
		return ~0ull;
	}
	return out;
}

} // namespace flshost
