// fls_fsst.cpp — FSST symbol-table trainer + greedy encoder (see fls_fsst.hpp).
#include "fls_fsst.hpp"

#include <algorithm>
#include <cstring>
#include <unordered_map>

namespace flshost {

namespace {

constexpr uint32_t kSlots = 1u << 12; // lookup slots (255 symbols: load factor 6 %)

inline uint64_t load_prefix(const uint8_t* s, uint32_t len) { // first `len` (<= 8) bytes, little endian, zero padded
	uint64_t v = 0;
	std::memcpy(&v, s, len);
	return v;
}

inline uint32_t hash_sym(uint64_t v, uint32_t len) {
	const uint64_t h = (v ^ ((uint64_t)len << 59)) * 0x9E3779B97F4A7C15ull;
	return (uint32_t)(h >> 40) & (kSlots - 1);
}

struct Cand {
	uint64_t value;
	uint32_t len;
	uint64_t gain;
};

struct KeyHash {
	size_t operator()(const std::pair<uint64_t, uint32_t>& k) const {
		return (size_t)((k.first ^ ((uint64_t)k.second << 59)) * 0x9E3779B97F4A7C15ull);
	}
};

} // namespace

void FsstEncoder::rebuild_lookup() {
	slots.assign(kSlots, 0);
	std::memset(byte_code, 0, sizeof(byte_code));
	max_len = 0;
	for (uint32_t c = 0; c < n_symbols; ++c) {
		const uint32_t len = table.len[c];
		max_len = std::max<uint8_t>(max_len, (uint8_t)len);
		if (len == 1) {
			byte_code[table.symbol[c] & 0xff] = (uint16_t)(c + 1);
			continue;
		}
		uint32_t h = hash_sym(table.symbol[c], len);
		while (slots[h]) {
			h = (h + 1) & (kSlots - 1);
		}
		slots[h] = c + 1;
	}
}

// longest symbol matching at s; returns its code or -1
int FsstEncoder::find(const uint8_t* s, uint64_t remaining, uint32_t* len_out) const {
	const uint32_t top = (uint32_t)std::min<uint64_t>(max_len, remaining);
	for (uint32_t len = top; len >= 2; --len) {
		const uint64_t v = load_prefix(s, len);
		uint32_t h = hash_sym(v, len);
		while (slots[h]) {
			const uint32_t c = slots[h] - 1;
			if (table.len[c] == len && table.symbol[c] == v) {
				*len_out = len;
				return (int)c;
			}
			h = (h + 1) & (kSlots - 1);
		}
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
	for (int round = 0; round < 5; ++round) {
		std::unordered_map<std::pair<uint64_t, uint32_t>, uint64_t, KeyHash> count;
		count.reserve(1 << 14);
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
				++count[{v, l}];
				if (prev_len && prev_len + l <= 8) { // candidate: the two adjacent symbols glued together
					++count[{prev_v | (v << (8 * prev_len)), prev_len + l}];
				}
				prev_v = v;
				prev_len = l;
				p += l;
			}
		}
		std::vector<Cand> cands;
		cands.reserve(count.size());
		for (const auto& kv : count) {
			// a single byte that is not a symbol costs 2 output bytes (escape), so a 1-byte symbol saves 1 per use like a
			// 2-byte symbol does; longer symbols save len - 1
			const uint64_t gain = kv.second * (kv.first.second == 1 ? 1 : kv.first.second - 1) + (kv.first.second == 1 ? kv.second / 2 : 0);
			cands.push_back({kv.first.first, kv.first.second, gain});
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
namespace {
constexpr uint32_t kSlots12 = 1u << 14;
inline uint32_t hash_sym12(uint64_t v, uint32_t len) {
	const uint64_t h = (v ^ ((uint64_t)len << 59)) * 0x9E3779B97F4A7C15ull;
	return (uint32_t)(h >> 40) & (kSlots12 - 1);
}
} // namespace

void Fsst12Encoder::rebuild_lookup() {
	slots.assign(kSlots12, 0);
	max_len = 0;
	for (uint32_t i = 0; i < n_symbols; ++i) {
		const uint32_t c = 256 + i, len = table.len[c];
		max_len = std::max<uint8_t>(max_len, (uint8_t)len);
		uint32_t h = hash_sym12(table.symbol[c], len);
		while (slots[h]) {
			h = (h + 1) & (kSlots12 - 1);
		}
		slots[h] = c;
	}
}

int Fsst12Encoder::find(const uint8_t* s, uint64_t remaining, uint32_t* len_out) const {
	const uint32_t top = (uint32_t)std::min<uint64_t>(max_len, remaining);
	for (uint32_t len = top; len >= 2; --len) {
		const uint64_t v = load_prefix(s, len);
		uint32_t h = hash_sym12(v, len);
		while (slots[h]) {
			const uint32_t c = slots[h];
			if (table.len[c] == len && table.symbol[c] == v) {
				*len_out = len;
				return (int)c;
			}
			h = (h + 1) & (kSlots12 - 1);
		}
	}
	return -1;
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
	for (int round = 0; round < 5; ++round) {
		std::unordered_map<std::pair<uint64_t, uint32_t>, uint64_t, KeyHash> count;
		count.reserve(1 << 16);
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
					++count[{v, l}];
				}
				if (prev_len && prev_len + l <= 8) {
					++count[{prev_v | (v << (8 * prev_len)), prev_len + l}];
				}
				prev_v = v;
				prev_len = l;
				p += l;
			}
		}
		std::vector<Cand> cands;
		cands.reserve(count.size());
		for (const auto& kv : count) {
			if (kv.second >= 2) { // a symbol seen once saves nothing worth a table slot
				cands.push_back({kv.first.first, kv.first.second, kv.second * (kv.first.second - 1)});
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
