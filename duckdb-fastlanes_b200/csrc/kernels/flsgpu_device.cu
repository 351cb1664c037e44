// flsgpu_device.cu — device side of the C ABI (include/flsgpu.h): staging of compressed row groups into HBM, launch
// planning for the decode kernel families, the pipelined host-to-host scan, pinned allocations.
//
// Replaces, for the read_fls scan path, what the reference does per worker thread in
// FastLanesReader::TryInitializeScan / DecodePhysicalColumns (reference src/reader/fls_reader.cpp:337-355,436-472):
// get_rowgroup_reader (-> stage), get_chunk + PointTo + KernelTraits<Op>::Decode (-> one fused kernel launch over
// every vector of the requested row groups).
#include <cuda_runtime.h>

#include <atomic>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "fls_error.hpp"
#include "fls_launch.hpp"
#include "fls_table.hpp"
#include "flsgpu.h"

namespace {

std::atomic<uint64_t> g_launches {0};
std::atomic<uint64_t> g_h2d_bytes {0}, g_d2h_bytes {0}; // bytes this process asked the copy engines to move (flsgpu_transfer_bytes)
std::mutex g_mu; // guards device selection + per-function attribute cache
// Devices of this process (flsgpu_init): CUDA ordinals, in the order given.  A table handle is bound to ONE of them
// (flsgpu_table_bind_device; index 0 unless told otherwise); one process per GPU passes a single ordinal, a single-process
// multi-GPU host (DuckDB: one process, N worker threads) opens one handle per device over the same image (flsgpu_table_share).
std::vector<int> g_devices;
std::map<int, int> g_n_sms_of;                                      // ordinal -> SM count
std::map<std::pair<const void*, int>, int> g_func_ready;            // (fn, ordinal) -> 1 once the smem opt-in is set
std::map<std::tuple<const void*, uint32_t, int>, int> g_occ;        // (fn, smem, ordinal) -> resident CTAs per SM

constexpr uint32_t kMaxSmemPerCta = 227u * 1024u;

// tuning knobs (environment overrides are for experiments; defaults are the measured best)
// A value that is not a plain decimal number inside [lo, hi] is reported once on stderr and ignored: a typo in an experiment
// knob must not turn into a zero-sized slab or a 4 GiB shared-memory request.
uint32_t env_u32(const char* name, uint32_t def, uint32_t lo = 0, uint32_t hi = 0xffffffffu) {
	const char* v = std::getenv(name);
	if (!v) {
		return def;
	}
	char* end = nullptr;
	const unsigned long long x = std::strtoull(v, &end, 10);
	if (end == v || *end != '\0' || x < lo || x > hi) {
		static std::mutex mu;
		static std::vector<std::string> reported;
		std::lock_guard<std::mutex> g(mu);
		if (std::find(reported.begin(), reported.end(), name) == reported.end()) {
			reported.push_back(name);
			std::fprintf(stderr, "flsgpu: ignoring %s=\"%s\" (expected an integer in [%u, %u]); using %u\n", name, v, lo, hi, def);
		}
		return def;
	}
	return (uint32_t)x;
}
const uint32_t g_run_shift = std::min(5u, env_u32("FLSGPU_RUN_SHIFT", 0, 0, 5));
const uint32_t g_prefetch = env_u32("FLSGPU_PREFETCH", 1, 0, 1);
const uint32_t g_pdl = env_u32("FLSGPU_PDL", 1, 0, 1);                // programmatic dependent launch between materialising decode launches
const uint32_t g_force_slots = env_u32("FLSGPU_NSLOTS", 0, 0, 16);     // experiments: pipeline depth override
const uint32_t g_max_ctas = env_u32("FLSGPU_MAX_CTAS", 0, 0, 32);      // experiments: cap on resident CTAs per SM
const uint32_t g_budget_kb = env_u32("FLSGPU_WARP_KB", 18, 4, 27);     // shared-memory budget per warp (KiB) the depth is sized for (14 -> 18 with chained launches: equal or up to +2.5 % across widths, all_types step 1.162 -> 1.149 ms)
// DICT launches: a warp takes runs of 8 consecutive vectors (so a CTA works through one row group = one dictionary) and
// the pipeline is kept shallow, which leaves most of the SM's L1 to the dictionary: 4096 x 8-byte keys 0.82 -> 0.26 ms
// per 128 M values (measured; low-cardinality dictionaries are unaffected)
const uint32_t g_dict_run_shift = std::min(5u, env_u32("FLSGPU_DICT_RUN_SHIFT", 3, 0, 5));
const uint32_t g_dict_kb = env_u32("FLSGPU_DICT_KB", 6, 2, 27);
const uint32_t g_kind_maps = env_u32("FLSGPU_KIND_MAPS", 1, 0, 1); // per-kind vector lists for columns that mix kernel families
const uint32_t g_swar8_kb = env_u32("FLSGPU_SWAR8_KB", 8, 2, 27);      // same for the 8-bit SWAR kernel (48 registers: up to five resident CTAs per SM)
// 8 / 16-bit columns decode more values per staged byte: resident warps matter more than pipeline depth (measured on 32 M
// SMALLINT values, 18 -> 8 KiB: FFOR w=16 26.5 -> 17.5 us, DELTA w=10/12 27 -> 18 us, w <= 8 unchanged or 12 % faster;
// profiles/r02_sweep_knobs.txt)
const uint32_t g_narrow_kb = env_u32("FLSGPU_NARROW_KB", 8, 2, 27);

int cuda_fail(cudaError_t e, const char* where) {
	char msg[256];
	std::snprintf(msg, sizeof(msg), "%s: %s", where, cudaGetErrorString(e));
	return fls_set_error(FLSGPU_ERR_CUDA, msg);
}
#define FLS_CUDA(call)                                                                                               \
	do {                                                                                                             \
		const cudaError_t e_ = (call);                                                                               \
		if (e_ != cudaSuccess) {                                                                                     \
			return cuda_fail(e_, #call);                                                                             \
		}                                                                                                            \
	} while (0)

// Make device `index` of the process's list current for the calling thread (index 0: the default device).
int ensure_device(int index = 0, int* ordinal_out = nullptr, int* n_sms_out = nullptr) {
	std::lock_guard<std::mutex> g(g_mu);
	if (g_devices.empty()) {
		int n = 0;
		const cudaError_t e = cudaGetDeviceCount(&n);
		if (e != cudaSuccess || n == 0) {
			return fls_set_error(FLSGPU_ERR_CUDA, "no CUDA device available (flsgpu has no CPU fallback)");
		}
		g_devices.push_back(0);
	}
	if (index < 0 || index >= (int)g_devices.size()) {
		return fls_set_error(FLSGPU_ERR_INVALID, "device index outside the list given to flsgpu_init");
	}
	const int ordinal = g_devices[index];
	FLS_CUDA(cudaSetDevice(ordinal));
	auto it = g_n_sms_of.find(ordinal);
	if (it == g_n_sms_of.end()) {
		int n_sms = 0;
		FLS_CUDA(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, ordinal));
		it = g_n_sms_of.emplace(ordinal, n_sms).first;
	}
	if (ordinal_out) {
		*ordinal_out = ordinal;
	}
	if (n_sms_out) {
		*n_sms_out = it->second;
	}
	return FLSGPU_OK;
}

// The device a raw device pointer / the caller's stream belongs to is the caller's business: entry points that take only
// device pointers (flsgpu_agg_reset, flsgpu_sync) run on the device that owns the pointer when it can be told, else on the
// thread's current device if it is one of ours, else on the default device.
int ensure_device_of(const void* dev_ptr) {
	if (dev_ptr) {
		cudaPointerAttributes attr {};
		if (cudaPointerGetAttributes(&attr, dev_ptr) == cudaSuccess && attr.type == cudaMemoryTypeDevice) {
			FLS_CUDA(cudaSetDevice(attr.device));
			return FLSGPU_OK;
		}
		cudaGetLastError();
	}
	return ensure_device();
}

} // namespace

// ---- per-table device state ---------------------------------------------------------------------------------------
struct FlsColumnPlan { // what a decode launch over (col, rg range) needs to know about the chunks it covers
	// 0: light (integer encodings), 1: ALP, 2: ALP_RD, 3: DELTA (T = 64), 4: FFOR + DELTA of 8-bit columns (SWAR kernel
	// when the output is 8 bits wide too, else folded back into 0 at launch time)
	bool has_main[5] = {false, false, false, false, false};
	bool has_rle = false, has_dict[3] = {false, false, false};
	bool main_needs_stage = false;
	uint32_t main_slot[5] = {0, 0, 0, 0, 0}, rle_slot = 0, dict_slot[3] = {0, 0, 0};
	// Columns whose chunks fall into several launch kinds (main families 0..4, DICT by code width 5..7, RLE 8): the vectors of
	// every kind, listed on the device, so that a kind's launch walks its own vectors only (a launch over the whole range spends
	// a descriptor fetch on every vector of the other kinds: the ALP_RD launch of the bench's DOUBLE column took 20 us to write
	// 0.07 MB).  nullptr: the launch walks the range.
	const uint32_t* kind_map[9] = {};
	uint32_t kind_n[9] = {};
};

struct FlsScanScratch { // two streams + two device slabs: H2D / decode / D2H of neighbouring slabs overlap
	cudaStream_t streams[2] = {nullptr, nullptr};
	void* out[2] = {nullptr, nullptr};
	uint64_t out_bytes = 0;
	// compact scans: the vector maps of a slab are put together in pinned memory (an asynchronous copy out of pageable
	// memory would first synchronise the stream, i.e. wait for the previous column's D2H copy)
	uint32_t* map_host[2] = {nullptr, nullptr};
	uint64_t map_bytes = 0;
	uint32_t* count_host = nullptr; // selected-rows scans: the slab's row count comes back here (pinned, one word per stream)
};

struct FlsDeviceState {
	int device = 0; // CUDA ordinal
	int n_sms = 0;
	uint8_t* d_image = nullptr; // bytes [0, footer_offset) of the container, same offsets as the file
	uint64_t image_bytes = 0;
	FlsChunkDev* d_chunks = nullptr; // [rg * n_columns + col]
	uint32_t* d_run_starts = nullptr; // CROSS_RLE: exclusive prefix sums of the run lengths, all chunks back to back
	uint64_t* d_heap_prefix = nullptr; // string columns: [rg * n_columns + col] = 16-byte padded decoded bytes of the column's earlier chunks
	std::vector<uint64_t> heap_prefix; // host copy (n_chunks + n_columns entries: the last row holds the column totals)
	uint64_t* d_key_prefix = nullptr;  // string columns: dictionary keys (DICT_STR / FSST_DICT aux_count) of the column's earlier chunks
	std::vector<uint64_t> key_prefix;
	std::vector<uint8_t> chunk_resident; // [rg * n_columns + col]: the chunk's H2D copy has been ENQUEUED (see chunk_stage)
	// Residency across streams: every staging pass records an event behind its copies (a ring of kStageEvents, a slot is
	// synchronised before it is reused, so a ticket older than the ring is known complete); chunk_stage[i] = ticket of the
	// pass that copied chunk i.  A consumer on another stream waits for the tickets of the chunks it reads (wait_resident).
	static constexpr uint32_t kStageEvents = 64;
	cudaEvent_t stage_events[kStageEvents] = {};
	cudaStream_t stage_streams[kStageEvents] = {};
	uint64_t stage_ticket = 0;           // tickets handed out so far (ticket k lives in ring slot k % kStageEvents)
	std::vector<uint64_t> chunk_stage;   // [rg * n_columns + col]: ticket + 1, 0 = none
	uint64_t resident_bytes = 0;
	bool host_pinned_tried = false;
	bool host_registered = false;
	std::map<std::tuple<uint32_t, uint32_t, uint32_t>, FlsColumnPlan> plans;
	std::vector<void*> plan_maps; // device blocks behind FlsColumnPlan::kind_map: freed with the table (launches may be in flight)
	uint64_t plan_map_bytes = 0;
	// pipelined scans (flsgpu_scan_columns_to_host): one scratch per concurrent caller, recycled through a free list
	std::vector<FlsScanScratch*> scratch_all, scratch_free;
	std::mutex mu;
};

void fls_device_release(flsgpu_table* t) {
	FlsDeviceState* d = t->dev;
	if (!d) {
		return;
	}
	cudaSetDevice(d->device);
	cudaDeviceSynchronize();
	for (FlsScanScratch* sc : d->scratch_all) {
		for (int i = 0; i < 2; ++i) {
			if (sc->streams[i]) {
				cudaStreamDestroy(sc->streams[i]);
			}
			cudaFree(sc->out[i]);
			cudaFreeHost(sc->map_host[i]);
		}
		cudaFreeHost(sc->count_host);
		delete sc;
	}
	for (void* p : d->plan_maps) {
		cudaFree(p);
	}
	cudaFree(d->d_image);
	cudaFree(d->d_chunks);
	cudaFree(d->d_run_starts);
	cudaFree(d->d_heap_prefix);
	cudaFree(d->d_key_prefix);
	for (cudaEvent_t e : d->stage_events) {
		if (e) {
			cudaEventDestroy(e);
		}
	}
	if (d->host_registered) {
		cudaHostUnregister(const_cast<uint8_t*>(t->bytes));
	}
	delete d;
	t->dev = nullptr;
}

namespace {

// Allocate the HBM arena + device chunk table on first use.  Pointers in the chunk table are final from the start
// (arena offsets equal file offsets); residency only says whether the bytes behind them have been copied.
int device_state_create(flsgpu_table* t, FlsDeviceState* d);

std::mutex g_create_mu; // serialises first-use creation: worker threads of one scan all arrive here at once

int device_state(flsgpu_table* t, FlsDeviceState** out) {
	std::lock_guard<std::mutex> g(g_create_mu);
	if (t->dev) {
		*out = t->dev;
		return cudaSetDevice(t->dev->device) == cudaSuccess ? FLSGPU_OK : fls_set_error(FLSGPU_ERR_CUDA, "cudaSetDevice failed");
	}
	int ordinal = 0, n_sms = 0;
	const int rc = ensure_device(t->device_index, &ordinal, &n_sms);
	if (rc) {
		return rc;
	}
	if (t->header.vec_per_rowgroup > 65535u) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "more than 65535 vectors per row group are not supported on the device");
	}
	auto d = new FlsDeviceState();
	d->device = ordinal;
	d->n_sms = n_sms;
	const int rc2 = device_state_create(t, d);
	if (rc2) { // publish only fully built states
		cudaFree(d->d_image);
		cudaFree(d->d_chunks);
		cudaFree(d->d_run_starts);
		cudaFree(d->d_heap_prefix);
		cudaFree(d->d_key_prefix);
		delete d;
		return rc2;
	}
	t->dev = d;
	*out = d;
	return FLSGPU_OK;
}

int device_state_create(flsgpu_table* t, FlsDeviceState* d) {
	const fls_file_header& h = t->header;
	const uint64_t n_chunks = (uint64_t)h.n_rowgroups * h.n_columns;
	d->image_bytes = h.footer_offset;
	d->chunk_resident.assign(n_chunks, 0);
	d->chunk_stage.assign(n_chunks, 0);
	FLS_CUDA(cudaMalloc(&d->d_image, d->image_bytes + 256));

	// CROSS_RLE run starts (reference cross_rle_kernel.hpp:8-21 rescans run lengths per vector; here the prefix sums
	// are built once so that a vector finds its first run by binary search)
	std::vector<uint64_t> starts_off(n_chunks, ~0ull);
	uint64_t n_starts = 0;
	for (uint64_t i = 0; i < n_chunks; ++i) {
		if (t->chunks[i].encoding == FLS_ENC_CROSS_RLE || t->chunks[i].encoding == FLS_ENC_CROSS_RLE_STR) {
			starts_off[i] = n_starts;
			n_starts += (uint64_t)t->chunks[i].aux_count + 1;
		}
	}
	if (n_starts) {
		std::vector<uint32_t> starts(n_starts);
		for (uint64_t i = 0; i < n_chunks; ++i) {
			if (starts_off[i] == ~0ull) {
				continue;
			}
			const fls_chunk_desc& c = t->chunks[i];
			const uint32_t* lens = reinterpret_cast<const uint32_t*>(t->bytes + c.offset + c.seg_off[FLS_SEG_AUX1]) +
			                       (c.encoding == FLS_ENC_CROSS_RLE_STR ? c.aux_count : 0u); // strings: the end offsets come first
			uint32_t acc = 0;
			uint32_t* s = starts.data() + starts_off[i];
			for (uint32_t r = 0; r < c.aux_count; ++r) {
				s[r] = acc;
				acc += lens[r];
			}
			s[c.aux_count] = acc;
		}
		FLS_CUDA(cudaMalloc(&d->d_run_starts, n_starts * sizeof(uint32_t)));
		FLS_CUDA(cudaMemcpy(d->d_run_starts, starts.data(), n_starts * sizeof(uint32_t), cudaMemcpyHostToDevice));
	}

	// string heap layout: a chunk's decoded bytes start at the padded prefix sum of the column's chunks before it
	bool any_string = false;
	for (uint32_t c = 0; c < h.n_columns; ++c) {
		any_string |= t->columns[c].phys_bits == 128;
	}
	if (any_string) {
		d->heap_prefix.assign(n_chunks + h.n_columns, 0);
		d->key_prefix.assign(n_chunks + h.n_columns, 0);
		for (uint32_t c = 0; c < h.n_columns; ++c) {
			uint64_t acc = 0, keys = 0;
			for (uint32_t r = 0; r <= h.n_rowgroups; ++r) {
				d->heap_prefix[(uint64_t)r * h.n_columns + c] = acc;
				d->key_prefix[(uint64_t)r * h.n_columns + c] = keys;
				if (r < h.n_rowgroups && t->columns[c].phys_bits == 128) {
					const fls_chunk_desc& ch = t->chunk(r, c);
					acc += (ch.const_val + 15u) & ~15ull;
					keys += FLS_ENC_IS_STR_DICT(ch.encoding) ? ch.aux_count : 0u;
				}
			}
		}
		FLS_CUDA(cudaMalloc(&d->d_heap_prefix, d->heap_prefix.size() * sizeof(uint64_t)));
		FLS_CUDA(cudaMemcpy(d->d_heap_prefix, d->heap_prefix.data(), d->heap_prefix.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
		FLS_CUDA(cudaMalloc(&d->d_key_prefix, d->key_prefix.size() * sizeof(uint64_t)));
		FLS_CUDA(cudaMemcpy(d->d_key_prefix, d->key_prefix.data(), d->key_prefix.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
	}

	std::vector<FlsChunkDev> host_chunks(n_chunks);
	for (uint64_t i = 0; i < n_chunks; ++i) {
		const fls_chunk_desc& c = t->chunks[i];
		FlsChunkDev& o = host_chunks[i];
		const uint8_t* base = d->d_image + c.offset;
		o.data = base + c.seg_off[FLS_SEG_DATA];
		o.meta = reinterpret_cast<const fls_vec_meta*>(base + c.seg_off[FLS_SEG_META]);
		o.aux0 = base + c.seg_off[FLS_SEG_AUX0];
		o.aux1 = c.encoding == FLS_ENC_CROSS_RLE ? reinterpret_cast<const uint8_t*>(d->d_run_starts + starts_off[i])
		                                         : base + c.seg_off[FLS_SEG_AUX1];
		if (c.encoding == FLS_ENC_CROSS_RLE_STR) { // no DATA segment: the slot carries the run starts, AUX1 keeps the end offsets
			o.data = reinterpret_cast<const uint8_t*>(d->d_run_starts + starts_off[i]);
		}
		o.const_val = c.const_val;
		o.aux_count = c.aux_count;
		o.n_vec = (uint16_t)c.n_vec;
		o.encoding = c.encoding;
		o.idx_bits = c.idx_bits;
		std::memcpy(o.rd_dict, c.rd_dict, sizeof(o.rd_dict));
	}
	FLS_CUDA(cudaMalloc(&d->d_chunks, n_chunks * sizeof(FlsChunkDev)));
	FLS_CUDA(cudaMemcpy(d->d_chunks, host_chunks.data(), n_chunks * sizeof(FlsChunkDev), cudaMemcpyHostToDevice));
	return FLSGPU_OK;
}

// Page-lock the container image once so that H2D staging runs at PCIe speed and asynchronously.  Failure is not an
// error (the image may already be pinned by flsgpu_host_alloc, or be a read-only mapping): copies then fall back to
// the driver's pageable path.
void pin_image(flsgpu_table* t, FlsDeviceState* d) {
	if (d->host_pinned_tried) {
		return;
	}
	d->host_pinned_tried = true;
	cudaPointerAttributes attr {};
	if (cudaPointerGetAttributes(&attr, t->bytes) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
		return; // already pinned
	}
	cudaGetLastError();
	if (cudaHostRegister(const_cast<uint8_t*>(t->bytes), t->size, cudaHostRegisterPortable) == cudaSuccess) {
		d->host_registered = true;
	} else {
		cudaGetLastError();
	}
}

// Copy the chunks of the listed columns (every column when cols == nullptr) of row groups [rg_begin, rg_begin+n_rg)
// that are not resident yet.  Chunks that are adjacent in the file - or separated by less than kMergeGap bytes of
// chunks nobody asked for - are merged into one cudaMemcpyAsync: a copy call costs a few microseconds of PCIe time,
// as much as ~256 KiB of payload.
constexpr uint64_t kMergeGap = 256u << 10;

int stage_range(flsgpu_table* t, FlsDeviceState* d, uint32_t rg_begin, uint32_t n_rg, const uint32_t* cols, uint32_t n_sel,
                cudaStream_t stream, const uint8_t* need = nullptr) {
	// need (optional): [(r - rg_begin) * n_columns + c] = 0 for chunks of the listed columns this pass may leave out
	// (compact scans: chunks without a FLAT vector are never read on the device)
	pin_image(t, d);
	const uint32_t n_cols = t->header.n_columns;
	std::vector<uint8_t> want(n_cols, cols ? 0 : 1);
	for (uint32_t i = 0; cols && i < n_sel; ++i) {
		want[cols[i]] = 1;
	}
	std::vector<uint64_t> staged; // chunks whose copy this pass enqueues: marked resident only once every copy is enqueued
	uint64_t run_begin = 0, run_end = 0, bytes = 0; // pending merged copy [begin, end) in file offsets
	auto flush = [&]() -> int {
		if (run_end > run_begin) {
			FLS_CUDA(cudaMemcpyAsync(d->d_image + run_begin, t->bytes + run_begin, run_end - run_begin, cudaMemcpyHostToDevice,
			                         stream));
			bytes += run_end - run_begin;
			g_h2d_bytes.fetch_add(run_end - run_begin, std::memory_order_relaxed);
		}
		run_begin = run_end = 0;
		return FLSGPU_OK;
	};
	int rc = FLSGPU_OK;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg && rc == FLSGPU_OK; ++r) {
		for (uint32_t c = 0; c < n_cols && rc == FLSGPU_OK; ++c) {
			const uint64_t i = (uint64_t)r * n_cols + c;
			if (!want[c] || d->chunk_resident[i] || (need && !need[(uint64_t)(r - rg_begin) * n_cols + c])) {
				continue;
			}
			staged.push_back(i);
			const fls_chunk_desc& ch = t->chunks[i];
			if (ch.is_equal) { // EXP_EQUAL: the bytes are those of another column's chunk; uploaded once
				const uint32_t src_col = ch.equal_col[0] | ((uint32_t)ch.equal_col[1] << 8);
				const uint64_t src = (uint64_t)r * n_cols + src_col;
				const bool src_in_pass = src < i && want[src_col] && (!need || need[(uint64_t)(r - rg_begin) * n_cols + src_col]);
				if (src < i && (d->chunk_resident[src] || src_in_pass)) {
					if (d->chunk_resident[src] && !src_in_pass) {
						staged.pop_back(); // copied by an earlier pass: inherit its ticket instead of this pass's
						d->chunk_stage[i] = d->chunk_stage[src];
						d->chunk_resident[i] = 1;
					}
					continue; // resident already, or uploaded earlier in this very pass (source columns come first)
				}
			}
			if (run_end > run_begin && ch.offset >= run_end && ch.offset - run_end <= kMergeGap) {
				run_end = ch.offset + ch.size;
			} else {
				rc = flush();
				run_begin = ch.offset;
				run_end = ch.offset + ch.size;
			}
		}
	}
	if (rc == FLSGPU_OK) {
		rc = flush();
	}
	if (rc != FLSGPU_OK || staged.empty()) {
		return rc; // nothing is marked: a failed pass leaves the chunks to the next one
	}
	// the event behind this pass's copies
	const uint32_t slot = (uint32_t)(d->stage_ticket % FlsDeviceState::kStageEvents);
	if (!d->stage_events[slot]) {
		FLS_CUDA(cudaEventCreateWithFlags(&d->stage_events[slot], cudaEventDisableTiming));
	} else {
		FLS_CUDA(cudaEventSynchronize(d->stage_events[slot])); // kStageEvents passes old: long complete
	}
	FLS_CUDA(cudaEventRecord(d->stage_events[slot], stream));
	d->stage_streams[slot] = stream;
	const uint64_t ticket = d->stage_ticket++;
	for (uint64_t i : staged) {
		d->chunk_resident[i] = 1;
		d->chunk_stage[i] = ticket + 1;
	}
	d->resident_bytes += bytes;
	return FLSGPU_OK;
}

uint32_t blob_bytes_max(const fls_chunk_desc& c) {
	return c.max_blob_bytes; // largest packed blob + exception tail of the chunk, recorded by the writer
}

const FlsColumnPlan& plan_for(flsgpu_table* t, FlsDeviceState* d, uint32_t col, uint32_t rg_begin, uint32_t n_rg) {
	const auto key = std::make_tuple(col, rg_begin, n_rg);
	auto it = d->plans.find(key);
	if (it != d->plans.end()) {
		return it->second;
	}
	FlsColumnPlan p;
	std::vector<uint8_t> kind_of(n_rg); // launch kind of every row group's chunk
	uint32_t kinds_seen = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		const fls_chunk_desc& c = t->chunk(r, col);
		const uint32_t bytes = blob_bytes_max(c);
		int kind;
		if (c.encoding == FLS_ENC_DICT) {
			const int k = c.idx_bits == 8 ? 0 : c.idx_bits == 16 ? 1 : 2;
			p.has_dict[k] = true;
			p.dict_slot[k] = std::max(p.dict_slot[k], bytes);
			kind = 5 + k;
		} else if (c.encoding == FLS_ENC_RLE) {
			p.has_rle = true;
			p.rle_slot = std::max(p.rle_slot, bytes);
			kind = 8;
		} else {
			const int f = c.encoding == FLS_ENC_ALP ? 1 : c.encoding == FLS_ENC_ALP_RD ? 2
			              : (c.encoding == FLS_ENC_DELTA && c.phys_bits == 64) ? 3
			              : (c.phys_bits == 8 && (c.encoding == FLS_ENC_FFOR || c.encoding == FLS_ENC_DELTA)) ? 4 : 0;
			p.has_main[f] = true;
			p.main_slot[f] = std::max(p.main_slot[f], bytes);
			p.main_needs_stage |= c.encoding == FLS_ENC_DELTA;
			kind = f;
		}
		kind_of[r - rg_begin] = (uint8_t)kind;
		kinds_seen |= 1u << kind;
	}
	// several kinds in the range: list every kind's vectors on the device (one block; bounded per table: past the bound the
	// launches simply walk the whole range as before)
	const uint64_t map_bytes = (uint64_t)n_rg * t->header.vec_per_rowgroup * 4u;
	if ((kinds_seen & (kinds_seen - 1u)) != 0 && g_kind_maps && d->plan_map_bytes + map_bytes <= (256ull << 20)) {
		std::vector<uint32_t> all;
		all.reserve((size_t)n_rg * t->header.vec_per_rowgroup);
		uint32_t first[9] = {};
		for (int k = 0; k < 9; ++k) {
			first[k] = (uint32_t)all.size();
			for (uint32_t r = 0; r < n_rg && ((kinds_seen >> k) & 1u); ++r) {
				if (kind_of[r] == k) {
					const uint32_t nv = t->chunk(rg_begin + r, col).n_vec;
					for (uint32_t v = 0; v < nv; ++v) {
						all.push_back(r * t->header.vec_per_rowgroup + v);
					}
				}
			}
			p.kind_n[k] = (uint32_t)all.size() - first[k];
		}
		void* block = nullptr;
		if (!all.empty() && cudaMalloc(&block, all.size() * 4u) == cudaSuccess &&
		    cudaMemcpy(block, all.data(), all.size() * 4u, cudaMemcpyHostToDevice) == cudaSuccess) {
			d->plan_maps.push_back(block);
			d->plan_map_bytes += all.size() * 4u;
			for (int k = 0; k < 9; ++k) {
				p.kind_map[k] = p.kind_n[k] ? static_cast<const uint32_t*>(block) + first[k] : nullptr;
			}
		} else {
			cudaGetLastError();
			cudaFree(block);
		}
	}
	if (d->plans.size() > 4096) {
		d->plans.clear(); // the device blocks of the dropped plans stay until the table closes: launches may still read them
	}
	return d->plans.emplace(key, p).first->second;
}

// Pick pipeline depth / grid for `fn` and launch it.
int launch_family(const FlsDeviceState* d, const void* fn, FlsLaunch L, uint32_t slot_bytes, uint32_t stage_bytes, cudaStream_t stream,
                  uint32_t tile_bytes = 0, uint32_t budget_kb = 0) {
	if (!fn) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "no kernel instantiation for this (lane width, value width, output width)");
	}
	L.slot_bytes = std::max<uint32_t>(128u, (slot_bytes + 127u) & ~127u);
	L.stage_bytes = stage_bytes;
	// shared memory: aim for two resident CTAs per SM (~104 KB each) unless a single slot pair needs more
	const uint32_t per_warp_budget = (budget_kb ? budget_kb : g_budget_kb) * 1024u; // budget_kb: kernels that want > 2 CTAs/SM
	const uint32_t fixed = FLS_DESC_BYTES_PER_WARP + L.stage_bytes + tile_bytes; // per-warp bytes besides the pipeline slots
	uint32_t n_slots = fixed < per_warp_budget ? (per_warp_budget - fixed) / L.slot_bytes : 0;
	// Two resident CTAs per SM matter more than pipeline depth (measured: 64-bit DELTA, 2 CTAs x 1 slot 0.139 ms vs
	// 1 CTA x 2 slots 0.154 ms for 100 M values): take the depth the budget allows, down to a single slot; only blobs of
	// a few hundred bytes, which decode much faster than a bulk copy lands, get at least ~4 KiB in flight per warp.
	n_slots = std::max(n_slots, std::min(8u, 4096u / L.slot_bytes));
	n_slots = std::max(1u, std::min(8u, n_slots));
	// L2 prefetch of the next batch (fls_decode_main.cuh): only while the batch ahead of ALL warps fits L2 comfortably
	// (~75 000 vectors x blob bytes; measured: helps 128 B blobs by 12 %, hurts 1.3 KiB blobs by 9 %)
	if (L.slot_bytes > 512u) {
		L.prefetch = 0;
	}
	if (g_force_slots) {
		n_slots = g_force_slots;
	}
	L.n_slots = n_slots;
	uint32_t wpc = kFlsWarpsPerCta, smem = 0;
	for (;; wpc >>= 1) { // wide vectors (8 KiB blobs + staging + aggregate tile) may need a smaller CTA to fit 227 KiB
		const uint32_t bar_bytes = (wpc * n_slots * 8u + 127u) & ~127u;
		smem = bar_bytes + wpc * (n_slots * L.slot_bytes + fixed);
		if (smem <= kMaxSmemPerCta) {
			break;
		}
		if (wpc == 1) {
			return fls_set_error(FLSGPU_ERR_INVALID, "vector blob too large for the shared-memory pipeline");
		}
	}
	int occ = 0;
	{
		std::lock_guard<std::mutex> g(g_mu);
		if (!g_func_ready.count(std::make_pair(fn, d->device))) { // the opt-in is per device
			FLS_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmemPerCta));
			g_func_ready[std::make_pair(fn, d->device)] = 1;
		}
		const auto key = std::make_tuple(fn, smem ^ (wpc << 24), d->device);
		auto it = g_occ.find(key);
		if (it == g_occ.end()) {
			FLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, (int)wpc * 32, smem));
			g_occ[key] = occ;
		} else {
			occ = it->second;
		}
	}
	if (occ < 1) {
		return fls_set_error(FLSGPU_ERR_CUDA, "decode kernel cannot be resident with the requested shared memory");
	}
	const uint64_t want = ((uint64_t)L.total_vecs + wpc - 1) / wpc;
	if (g_max_ctas) {
		occ = std::min<int>(occ, (int)g_max_ctas);
	}
	const uint32_t grid = (uint32_t)std::min<uint64_t>(want, (uint64_t)occ * d->n_sms);
	// materialising launches (nothing but staged bytes read, nothing but the caller's column written): programmatic
	// stream serialization, the kernel releases its successor at once and waits for its predecessor before it exits
	L.pdl = (g_pdl && L.agg == nullptr && L.mask_mode == 0u) ? 1u : 0u;
	void* args[] = {&L};
	if (L.pdl) {
		cudaLaunchConfig_t cfg {};
		cfg.gridDim = dim3(grid);
		cfg.blockDim = dim3(wpc * 32);
		cfg.dynamicSmemBytes = smem;
		cfg.stream = stream;
		cudaLaunchAttribute attr {};
		attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
		attr.val.programmaticStreamSerializationAllowed = 1;
		cfg.attrs = &attr;
		cfg.numAttrs = 1;
		FLS_CUDA(cudaLaunchKernelExC(&cfg, fn, args));
	} else {
		FLS_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(wpc * 32), args, smem, stream));
	}
	g_launches.fetch_add(1, std::memory_order_relaxed);
	return FLSGPU_OK;
}

int check_filters(const flsgpu_table* t, const flsgpu_filter* f, uint32_t n, const char* who) {
	if (!t || (!f && n)) {
		return fls_set_error(FLSGPU_ERR_INVALID, "filter list is null");
	}
	for (uint32_t i = 0; i < n; ++i) {
		if (f[i].col >= t->header.n_columns || f[i].op < FLSGPU_CMP_EQ || f[i].op > FLSGPU_CMP_GE) {
			char msg[160];
			std::snprintf(msg, sizeof(msg), "%s: filter %u has a bad column or comparison", who, i);
			return fls_set_error(FLSGPU_ERR_INVALID, msg);
		}
		if (t->columns[f[i].col].phys_bits == 128 && !f[i].str_data && f[i].str_len) {
			return fls_set_error(FLSGPU_ERR_INVALID, "filter on a string column without the constant's bytes (str_data)");
		}
		if (t->columns[f[i].col].phys_bits == 128 && f[i].str_len > 0xffffffffull) {
			return fls_set_error(FLSGPU_ERR_INVALID, "string filter constant longer than a string_t can be");
		}
	}
	return FLSGPU_OK;
}


struct OutSpec {
	int ob = 0;        // bytes per output value
	uint32_t conv = 0; // 0 zero-extend / same width, 1 sign-extend, 2 f32 -> f64
};

// Destination type rules of the materializer (reference fls_type_resolver.hpp:22-33,120-131: same physical type ->
// copy, otherwise static_cast from the source viewed with the column's signedness, unffor_kernel.hpp:16-25).
int resolve_out(const fls_column_desc& col, int32_t out_type, OutSpec* o) {
	const int T = col.phys_bits;
	if (out_type == FLSGPU_OUT_NATIVE) {
		o->ob = T / 8;
		o->conv = 0;
		return FLSGPU_OK;
	}
	int ob = 0;
	bool out_float = false;
	switch (out_type) {
	case FLSGPU_OUT_I8:
	case FLSGPU_OUT_U8: ob = 1; break;
	case FLSGPU_OUT_I16:
	case FLSGPU_OUT_U16: ob = 2; break;
	case FLSGPU_OUT_I32:
	case FLSGPU_OUT_U32: ob = 4; break;
	case FLSGPU_OUT_I64:
	case FLSGPU_OUT_U64: ob = 8; break;
	case FLSGPU_OUT_F32: ob = 4; out_float = true; break;
	case FLSGPU_OUT_F64: ob = 8; out_float = true; break;
	default: return fls_set_error(FLSGPU_ERR_INVALID, "unknown flsgpu_out_type");
	}
	if ((bool)col.is_float != out_float) {
		return fls_set_error(FLSGPU_ERR_CONVERSION, "integer <-> floating point materialisation is not supported");
	}
	if (ob * 8 < T) {
		return fls_set_error(FLSGPU_ERR_CONVERSION, "destination type is narrower than the column's physical type");
	}
	o->ob = ob;
	o->conv = ob * 8 == T ? 0u : col.is_float ? 2u : col.is_signed ? 1u : 0u;
	return FLSGPU_OK;
}

struct MaskSpec { // filter pushdown on aggregate-family launches (FlsLaunch.mask_mode)
	uint32_t mode = 0, op = 0;
	uint64_t constant = 0;
	uint32_t op2 = 0xff; // second comparison on the same column, evaluated in the same pass
	uint64_t constant2 = 0;
	uint32_t* mask = nullptr;
	const uint8_t* rg_keep = nullptr;
};

int decode_column_impl(flsgpu_table* t, FlsDeviceState* d, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const OutSpec& os,
                       void* out_dev, cudaStream_t stream, flsgpu_agg_acc* agg = nullptr, const MaskSpec* ms = nullptr,
                       const uint32_t* vec_map_dev = nullptr, uint32_t n_mapped = 0) {
	const fls_file_header& h = t->header;
	const int T = t->columns[col].phys_bits;
	if (T == 128) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "string column: use flsgpu_decode_strings / flsgpu_decode_string_codes");
	}
	FlsColumnPlan p;
	{
		std::lock_guard<std::mutex> g(d->mu);
		p = plan_for(t, d, col, rg_begin, n_rg);
	}
	FlsLaunch L {};
	L.chunks = d->d_chunks + (uint64_t)rg_begin * h.n_columns + col;
	L.chunk_stride = h.n_columns;
	L.vec_per_rg = h.vec_per_rowgroup;
	L.total_vecs = vec_map_dev ? n_mapped : n_rg * h.vec_per_rowgroup;
	L.vec_map = vec_map_dev;
	L.conv = os.conv;
	L.out = out_dev;
	L.run_shift = g_run_shift;
	// fused aggregates write nothing (no read/write turnarounds to avoid); 8/16-bit outputs are instruction bound, not
	// bandwidth bound, and the second descriptor fetch costs them ~4 %
	L.prefetch = (agg || ms || os.ob < 4) ? 0u : g_prefetch;
	L.agg = agg;
	L.agg_signed = t->columns[col].is_signed;
	L.agg_float = t->columns[col].is_float;
	if (ms) {
		L.mask_mode = ms->mode;
		L.cmp_op = ms->op;
		L.cmp_const = ms->constant;
		L.cmp_op2 = ms->op2;
		L.cmp_const2 = ms->constant2;
		L.mask = ms->mask;
		L.rg_keep = ms->rg_keep;
	}
	const bool agg_family = agg || ms;
	const uint32_t tile = agg_family ? 128u * T : 0u; // aggregate launches decode into a per-warp shared-memory tile
	int rc = FLSGPU_OK;
	const uint32_t light = FLS_FAMILY_MAIN & ~((1u << FLS_ENC_ALP) | (1u << FLS_ENC_ALP_RD));
	const uint32_t swar8 = (1u << FLS_ENC_FFOR) | (1u << FLS_ENC_DELTA);
	const bool use_swar8 = p.has_main[4] && !agg_family && T == 8 && os.ob == 1; // else family 4's chunks decode in the light kernel
	const uint32_t fam_mask[5] = {T == 64 ? light & ~(1u << FLS_ENC_DELTA) : use_swar8 ? light & ~swar8 : light, 1u << FLS_ENC_ALP,
	                              1u << FLS_ENC_ALP_RD, 1u << FLS_ENC_DELTA, swar8};
	// FLOAT columns that mix ALP and ALP_RD chunks: one launch of the instantiation that carries both (same 128-register
	// budget as the ALP_RD kernel alone) instead of two launches with their ramp and tail
	const bool fuse_alp = T == 32 && !agg_family && p.has_main[1] && p.has_main[2];
	// a column that mixes launch kinds: every launch walks the list of its own vectors and decodes them in place
	const uint32_t range_vecs = L.total_vecs;
	const bool use_maps = !vec_map_dev && !agg_family;
	auto walk = [&](int kind) {
		const bool mapped = use_maps && kind >= 0 && p.kind_map[kind] != nullptr;
		L.vec_map = mapped ? p.kind_map[kind] : vec_map_dev;
		L.total_vecs = mapped ? p.kind_n[kind] : range_vecs;
		L.map_in_place = mapped ? 1u : 0u;
	};
	if (fuse_alp) {
		walk(-1);
		L.family_mask = fam_mask[1] | fam_mask[2];
		L.idx_bits = 0;
		rc = launch_family(d, fls_kernel_main(T, os.ob, 5), L, std::max(p.main_slot[1], p.main_slot[2]), 0u, stream, tile);
	}
	for (int f = 0; f < 5 && rc == FLSGPU_OK; ++f) {
		if (fuse_alp && (f == 1 || f == 2)) {
			continue;
		}
		const bool fold4 = f == 0 && p.has_main[4] && !use_swar8;
		if ((f == 4 ? use_swar8 : p.has_main[f]) || fold4) {
			walk(fold4 ? -1 : f); // a launch that also takes the 8-bit family's chunks walks the range
			L.family_mask = fam_mask[f];
			L.idx_bits = 0;
			// DELTA needs the per-warp transposition area (+ 16 lane totals for T = 64); the SWAR kernel stores from registers
			const bool stage = p.main_needs_stage && (f == 3 || (f == 0 && T != 64));
			const uint32_t slot = fold4 ? std::max(p.main_slot[0], p.main_slot[4]) : p.main_slot[f];
			rc = launch_family(d, agg_family ? fls_kernel_agg(0, T, T, f == 3 ? 0 : f) : fls_kernel_main(T, os.ob, f), L, slot,
			                   stage ? 128u * T + (T == 64 ? 128u : 0u) : 0u, stream, tile,
			                   f == 4 ? g_swar8_kb : (T <= 16 && !agg_family) ? g_narrow_kb : 0u);
		}
	}
	for (int k = 0; k < 3 && rc == FLSGPU_OK; ++k) {
		if (p.has_dict[k]) {
			const int it = 8 << k;
			walk(5 + k);
			L.family_mask = FLS_FAMILY_DICT;
			L.idx_bits = (uint32_t)it;
			L.run_shift = g_dict_run_shift;
			rc = launch_family(d, agg_family ? fls_kernel_agg(1, it, T, 0) : fls_kernel_dict(it, T, os.ob), L, p.dict_slot[k], 0, stream, tile,
			                   g_dict_kb + tile / 1024u);
		}
	}
	L.run_shift = g_run_shift;
	if (p.has_rle && rc == FLSGPU_OK) {
		walk(8);
		L.family_mask = FLS_FAMILY_RLE;
		L.idx_bits = 0;
		rc = launch_family(d, agg_family ? fls_kernel_agg(2, 16, T, 0) : fls_kernel_rle(T, os.ob), L, p.rle_slot, 2048u, stream, tile);
	}
	return rc;
}

int check_range(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const char* who) {
	if (!t || col >= t->header.n_columns || n_rg == 0 || rg_begin >= t->header.n_rowgroups ||
	    n_rg > t->header.n_rowgroups - rg_begin) {
		char msg[160];
		std::snprintf(msg, sizeof(msg), "%s: column or row-group range out of bounds", who);
		return fls_set_error(FLSGPU_ERR_INVALID, msg);
	}
	return FLSGPU_OK;
}

// Every chunk of the column in the range must have been staged; `stream` is made to wait for staging passes that ran on
// other streams and may still be in flight (flsgpu_table_stage on stream A followed by a decode on stream B; two
// concurrent host-to-host scans over the same row groups).
int check_resident(const flsgpu_table* t, FlsDeviceState* d, uint32_t col, uint32_t rg_begin, uint32_t n_rg, cudaStream_t stream) {
	std::lock_guard<std::mutex> g(d->mu);
	uint64_t waited = 0; // bit per ring slot already waited for in this call
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		const uint64_t i = (uint64_t)r * t->header.n_columns + col;
		if (!d->chunk_resident[i]) {
			return fls_set_error(FLSGPU_ERR_INVALID, "row group not staged: call flsgpu_table_stage first");
		}
		const uint64_t tk = d->chunk_stage[i];
		if (tk == 0 || d->stage_ticket - (tk - 1) > FlsDeviceState::kStageEvents) {
			continue; // the ring slot was reused since: that pass was synchronised
		}
		const uint32_t slot = (uint32_t)((tk - 1) % FlsDeviceState::kStageEvents);
		if (d->stage_streams[slot] == stream || ((waited >> slot) & 1ull)) {
			continue; // same stream: ordered already
		}
		waited |= 1ull << slot;
		FLS_CUDA(cudaStreamWaitEvent(stream, d->stage_events[slot], 0));
	}
	return FLSGPU_OK;
}

} // namespace

// ---- C ABI ---------------------------------------------------------------------------------------------------------
extern "C" {

int flsgpu_init(int n_devices, const int* device_ids) {
	int n = 0;
	const cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) {
		return fls_set_error(FLSGPU_ERR_CUDA, "flsgpu_init: no CUDA device available (flsgpu has no CPU fallback)");
	}
	std::vector<int> devs;
	for (int i = 0; i < n_devices && device_ids; ++i) {
		if (device_ids[i] < 0 || device_ids[i] >= n) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_init: device id out of range");
		}
		devs.push_back(device_ids[i]); // an ordinal may be listed more than once (several handles of a table on one device)
	}
	if (devs.empty()) {
		devs.push_back(0);
	}
	{
		std::lock_guard<std::mutex> g(g_mu);
		g_devices = devs;
	}
	return ensure_device();
}

int flsgpu_n_devices(int* out) {
	if (!out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_n_devices: null argument");
	}
	std::lock_guard<std::mutex> g(g_mu);
	*out = g_devices.empty() ? 1 : (int)g_devices.size();
	return FLSGPU_OK;
}

int flsgpu_table_bind_device(flsgpu_table* t, int device_index) {
	if (!t) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_bind_device: null table");
	}
	if (t->dev) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_bind_device: the table already has device state (bind before the first staging)");
	}
	{
		std::lock_guard<std::mutex> g(g_mu);
		if (device_index < 0 || device_index >= (int)std::max<size_t>(1, g_devices.size())) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_bind_device: index outside the list given to flsgpu_init");
		}
	}
	t->device_index = device_index;
	return FLSGPU_OK;
}

int flsgpu_device_count(int* out) {
	if (!out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_device_count: null argument");
	}
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) {
		cudaGetLastError();
		n = 0;
	}
	*out = n;
	return FLSGPU_OK;
}

int flsgpu_table_stage(flsgpu_table* t, uint32_t rg_begin, uint32_t n_rg, void* cuda_stream) {
	int rc = check_range(t, 0, rg_begin, n_rg, "flsgpu_table_stage");
	if (rc) {
		return rc;
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	std::lock_guard<std::mutex> g(d->mu);
	return stage_range(t, d, rg_begin, n_rg, nullptr, 0, static_cast<cudaStream_t>(cuda_stream));
}

int flsgpu_table_evict(flsgpu_table* t) {
	if (!t) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_evict: null table");
	}
	if (t->dev) {
		FLS_CUDA(cudaSetDevice(t->dev->device));
		FLS_CUDA(cudaDeviceSynchronize());
		std::lock_guard<std::mutex> g(t->dev->mu);
		std::fill(t->dev->chunk_resident.begin(), t->dev->chunk_resident.end(), 0);
		t->dev->resident_bytes = 0;
	}
	return FLSGPU_OK;
}

int flsgpu_table_unstage(flsgpu_table* t) {
	if (!t) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_unstage: null table");
	}
	fls_device_release(t);
	return FLSGPU_OK;
}

int flsgpu_table_resident_bytes(const flsgpu_table* t, uint64_t* out) {
	if (!t || !out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_table_resident_bytes: null argument");
	}
	*out = t->dev ? t->dev->resident_bytes : 0;
	return FLSGPU_OK;
}

int flsgpu_decode_column(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, int32_t out_type, void* out_dev,
                         uint64_t out_capacity_bytes, void* cuda_stream) {
	int rc = check_range(t, col, rg_begin, n_rg, "flsgpu_decode_column");
	if (rc) {
		return rc;
	}
	OutSpec os;
	rc = resolve_out(t->columns[col], out_type, &os);
	if (rc) {
		return rc;
	}
	const uint64_t need = (uint64_t)n_rg * t->header.vec_per_rowgroup * FLS_VEC_SZ * os.ob;
	if (!out_dev || (reinterpret_cast<uintptr_t>(out_dev) & 15u) || out_capacity_bytes < need) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_decode_column: output buffer null, misaligned or too small");
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	rc = check_resident(t, d, col, rg_begin, n_rg, static_cast<cudaStream_t>(cuda_stream));
	if (rc) {
		return rc;
	}
	return decode_column_impl(t, d, col, rg_begin, n_rg, os, out_dev, static_cast<cudaStream_t>(cuda_stream));
}

int flsgpu_rowgroup_decode(flsgpu_table* t, uint32_t rg, const uint32_t* col_ids, uint32_t n_cols, void* const* out_ptrs,
                           const int32_t* out_types, void* cuda_stream) {
	if (!t || !col_ids || !out_ptrs || n_cols == 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_rowgroup_decode: null argument");
	}
	for (uint32_t i = 0; i < n_cols; ++i) {
		int rc = check_range(t, col_ids[i], rg, 1, "flsgpu_rowgroup_decode");
		if (rc) {
			return rc;
		}
		OutSpec os;
		rc = resolve_out(t->columns[col_ids[i]], out_types ? out_types[i] : FLSGPU_OUT_NATIVE, &os);
		if (rc) {
			return rc;
		}
		if (!out_ptrs[i] || (reinterpret_cast<uintptr_t>(out_ptrs[i]) & 15u)) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_rowgroup_decode: output pointer null or not 16-byte aligned");
		}
	}
	FlsDeviceState* d = nullptr;
	int rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	for (uint32_t i = 0; i < n_cols; ++i) {
		rc = check_resident(t, d, col_ids[i], rg, 1, static_cast<cudaStream_t>(cuda_stream));
		if (rc) {
			return rc;
		}
		OutSpec os;
		resolve_out(t->columns[col_ids[i]], out_types ? out_types[i] : FLSGPU_OUT_NATIVE, &os);
		rc = decode_column_impl(t, d, col_ids[i], rg, 1, os, out_ptrs[i], static_cast<cudaStream_t>(cuda_stream));
		if (rc) {
			return rc;
		}
	}
	return FLSGPU_OK;
}

// Pipelined host-to-host scan of a projection (numeric columns, flat strings, dictionary-form strings), optionally with
// the selection mask of a filter set evaluated on the device slab by slab.
int flsgpu_scan_projection_to_host(flsgpu_table* t, const flsgpu_scan_column* cols, uint32_t n_cols, uint32_t rg_begin, uint32_t n_rg) {
	return flsgpu_scan_filtered_to_host(t, cols, n_cols, nullptr, 0, rg_begin, n_rg, nullptr, 0);
}

int flsgpu_scan_filtered_to_host(flsgpu_table* t, const flsgpu_scan_column* cols, uint32_t n_cols, const flsgpu_filter* filters,
                                 uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg, uint32_t* mask_host, uint64_t mask_capacity_bytes) {
	if (!t || !cols || n_cols == 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_projection_to_host: null argument");
	}
	if (n_filters) {
		const int frc = check_filters(t, filters, n_filters, "flsgpu_scan_filtered_to_host");
		if (frc) {
			return frc;
		}
		if (!mask_host || mask_capacity_bytes < (uint64_t)n_rg * t->header.vec_per_rowgroup * 128u) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_filtered_to_host: mask buffer null or too small");
		}
	}
	const fls_file_header& h = t->header;
	struct ColPlan {
		OutSpec os;
		bool is_string = false, dict = false, compact = false;
		uint64_t rg_bytes = 0;            // row bytes per row group (values, string_t or selection entries)
		uint64_t dev_rows = 0, dev_heap = 0, dev_keys = 0, dev_map = 0; // offsets inside a slab's device scratch
		std::vector<flsgpu_vector_form> own_forms; // compact: form of every vector of the range, unless the caller handed them in
		const flsgpu_vector_form* forms = nullptr;
	};
	std::vector<std::vector<uint32_t>> vec_maps; // compact: the FLAT vectors of a (slab, column), alive until the streams drain
	std::vector<ColPlan> cp(n_cols);
	uint64_t rg_bytes_all = 0;
	for (uint32_t i = 0; i < n_cols; ++i) {
		const flsgpu_scan_column& c = cols[i];
		int rc = check_range(t, c.col, rg_begin, n_rg, "flsgpu_scan_projection_to_host");
		if (rc) {
			return rc;
		}
		ColPlan& p = cp[i];
		p.is_string = t->columns[c.col].phys_bits == 128;
		if (p.is_string) {
			p.dict = c.string_form == FLSGPU_STRING_DICTIONARY;
			p.rg_bytes = (uint64_t)h.vec_per_rowgroup * FLS_VEC_SZ * (p.dict ? 4u : 16u);
			uint64_t heap_need = 0, keys_need = 0;
			flsgpu_string_heap_bytes(t, c.col, rg_begin, n_rg, &heap_need);
			if (p.dict) {
				rc = flsgpu_string_dict_keys(t, c.col, rg_begin, n_rg, &keys_need);
				if (rc) {
					return rc;
				}
			}
			if (!c.heap_host || c.heap_capacity_bytes < heap_need || (p.dict && (!c.keys_host || c.keys_capacity_bytes < keys_need * 16u))) {
				return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_projection_to_host: string heap / key buffer null or too small");
			}
		} else {
			rc = resolve_out(t->columns[c.col], c.out_type, &p.os);
			if (rc) {
				return rc;
			}
			p.rg_bytes = (uint64_t)h.vec_per_rowgroup * FLS_VEC_SZ * p.os.ob;
			p.compact = c.compact != 0;
		}
		rg_bytes_all += p.rg_bytes;
		uint64_t need_bytes = p.rg_bytes * n_rg;
		if (p.compact) { // only the FLAT vectors of the range cross the link, packed
			uint64_t n_flat = 0;
			const size_t n_vec_all = (size_t)n_rg * h.vec_per_rowgroup;
			if (c.forms) {
				p.forms = c.forms;
				for (size_t v = n_vec_all; v-- > 0;) { // slots are ranks: the last FLAT vector tells how many there are
					if (c.forms[v].form == FLS_VEC_FORM_FLAT) {
						n_flat = (uint64_t)c.forms[v].flat_slot + 1;
						break;
					}
				}
			} else {
				p.own_forms.resize(n_vec_all);
				rc = flsgpu_column_vector_forms(t, c.col, rg_begin, n_rg, p.own_forms.data(), &n_flat);
				if (rc) {
					return rc;
				}
				p.forms = p.own_forms.data();
			}
			need_bytes = n_flat * FLS_VEC_SZ * p.os.ob;
		}
		if ((!c.out_host && need_bytes) || c.out_capacity_bytes < need_bytes) {
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_projection_to_host: output buffer null or too small");
		}
	}
	FlsDeviceState* d = nullptr;
	int rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	auto heap_off = [&](uint32_t col, uint32_t r) { // bytes of heap before row group rg_begin + r, relative to the range
		const uint64_t i0 = (uint64_t)rg_begin * h.n_columns + col;
		return d->heap_prefix[i0 + (uint64_t)r * h.n_columns] - d->heap_prefix[i0];
	};
	auto key_off = [&](uint32_t col, uint32_t r) {
		const uint64_t i0 = (uint64_t)rg_begin * h.n_columns + col;
		return d->key_prefix[i0 + (uint64_t)r * h.n_columns] - d->key_prefix[i0];
	};
	// slabs of ~256 MiB decoded (all projected columns together): deep enough to hide launch latency and the per-copy
	// overhead (measured e2e, 100 M-row all_types: 16 MiB 42.1, 64 MiB 50.8, 256 MiB 52.1, 1 GiB 52.7 GB/s against a
	// 57 GB/s D2H link), small enough that H2D, decode and D2H of neighbouring slabs overlap on the two streams
	uint64_t rg_avg = rg_bytes_all;
	for (uint32_t i = 0; i < n_cols; ++i) {
		if (cp[i].is_string) {
			rg_avg += heap_off(cols[i].col, n_rg) / n_rg;
		}
	}
	const uint64_t slab_target = (uint64_t)env_u32("FLSGPU_SLAB_MB", 256, 1, 65536) << 20;
	const uint32_t slab_rg = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(n_rg, slab_target / std::max<uint64_t>(1, rg_avg)));
	uint64_t slab_bytes = 0;
	for (uint32_t i = 0; i < n_cols; ++i) {
		ColPlan& p = cp[i];
		p.dev_rows = slab_bytes;
		slab_bytes += p.rg_bytes * slab_rg;
		if (p.is_string) {
			uint64_t heap_max = 0, keys_max = 0;
			for (uint32_t r = 0; r < n_rg; r += slab_rg) {
				const uint32_t e = std::min(n_rg, r + slab_rg);
				heap_max = std::max(heap_max, heap_off(cols[i].col, e) - heap_off(cols[i].col, r));
				keys_max = std::max(keys_max, key_off(cols[i].col, e) - key_off(cols[i].col, r));
			}
			p.dev_heap = slab_bytes;
			slab_bytes += heap_max + 16;
			p.dev_keys = slab_bytes;
			slab_bytes += p.dict ? keys_max * 16u + 16u : 0u;
		}
	}
	for (uint32_t i = 0; i < n_cols; ++i) {
		if (cp[i].compact) {
			cp[i].dev_map = slab_bytes;
			slab_bytes += (((uint64_t)slab_rg * h.vec_per_rowgroup * 4u) + 255u) & ~255ull;
		}
	}
	const uint64_t dev_mask = slab_bytes, rg_mask_bytes = (uint64_t)h.vec_per_rowgroup * 128u;
	slab_bytes += n_filters ? rg_mask_bytes * slab_rg : 0u;
	FlsScanScratch* sc = nullptr;
	{
		std::lock_guard<std::mutex> g(d->mu);
		if (!d->scratch_free.empty()) {
			sc = d->scratch_free.back();
			d->scratch_free.pop_back();
		} else {
			sc = new FlsScanScratch();
			d->scratch_all.push_back(sc);
		}
	}
	struct Return { // hand the scratch back on every exit path
		FlsDeviceState* d;
		FlsScanScratch* sc;
		~Return() {
			std::lock_guard<std::mutex> g(d->mu);
			d->scratch_free.push_back(sc);
		}
	} ret {d, sc};
	if (sc->out_bytes < slab_bytes) {
		sc->out_bytes = 0; // a failed regrow must not leave the old size behind a freed or missing slab
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaFree(sc->out[i]));
			sc->out[i] = nullptr;
		}
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaMalloc(&sc->out[i], slab_bytes));
		}
		sc->out_bytes = slab_bytes;
	}
	for (int i = 0; i < 2; ++i) {
		if (!sc->streams[i]) {
			FLS_CUDA(cudaStreamCreateWithFlags(&sc->streams[i], cudaStreamNonBlocking));
		}
	}
	uint64_t n_compact = 0;
	for (uint32_t i = 0; i < n_cols; ++i) {
		n_compact += cp[i].compact ? 1 : 0;
	}
	const uint64_t map_need = n_compact * slab_rg * h.vec_per_rowgroup * 4u;
	if (sc->map_bytes < map_need) {
		sc->map_bytes = 0;
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaFreeHost(sc->map_host[i]));
			sc->map_host[i] = nullptr;
		}
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&sc->map_host[i]), map_need, cudaHostAllocDefault));
		}
		sc->map_bytes = map_need;
	}
	bool slab_used[2] = {false, false};
	std::vector<uint32_t> col_ids(n_cols);
	for (uint32_t i = 0; i < n_cols; ++i) {
		col_ids[i] = cols[i].col;
	}
	for (uint32_t i = 0; i < n_filters; ++i) { // filter columns are staged too, whether projected or not
		if (std::find(col_ids.begin(), col_ids.end(), filters[i].col) == col_ids.end()) {
			col_ids.push_back(filters[i].col);
		}
	}
	uint32_t slab = 0;
	int status = FLSGPU_OK;
	auto d2h = [&](void* dst, const void* src, uint64_t bytes, cudaStream_t s) {
		if (status == FLSGPU_OK && bytes && cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s) != cudaSuccess) {
			status = cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(D2H)");
		}
		g_d2h_bytes.fetch_add(bytes, std::memory_order_relaxed);
	};
	for (uint32_t r = 0; r < n_rg && status == FLSGPU_OK; r += slab_rg, ++slab) {
		const uint32_t n = std::min(slab_rg, n_rg - r);
		cudaStream_t s = sc->streams[slab & 1];
		if (slab_used[slab & 1] && n_compact) { // this slab's pinned map area was last read by the copies of slab - 2
			const cudaError_t e = cudaStreamSynchronize(s);
			if (e != cudaSuccess) {
				status = cuda_fail(e, "cudaStreamSynchronize");
				break;
			}
		}
		slab_used[slab & 1] = true;
		uint64_t map_used = 0; // entries of sc->map_host[slab & 1] taken so far
		// compact columns: the FLAT vectors of the slab (launch order = slot order) and the chunks that hold any
		std::vector<uint8_t> need;
		std::vector<uint32_t> map_of(n_cols, 0xffffffffu); // column -> index into vec_maps
		for (uint32_t i = 0; i < n_cols; ++i) {
			if (!cp[i].compact) {
				continue;
			}
			if (need.empty()) {
				need.assign((size_t)n * h.n_columns, 1);
			}
			vec_maps.emplace_back();
			map_of[i] = (uint32_t)vec_maps.size() - 1;
			std::vector<uint32_t>& vm = vec_maps.back();
			bool filtered = false;
			for (uint32_t k = 0; k < n_filters; ++k) {
				filtered |= filters[k].col == cols[i].col;
			}
			for (uint32_t rr = 0; rr < n; ++rr) {
				const size_t before = vm.size();
				for (uint32_t v = 0; v < h.vec_per_rowgroup; ++v) {
					if (cp[i].forms[(size_t)(r + rr) * h.vec_per_rowgroup + v].form == FLS_VEC_FORM_FLAT) {
						vm.push_back(rr * h.vec_per_rowgroup + v);
					}
				}
				if (vm.size() == before && !filtered) {
					need[(size_t)rr * h.n_columns + cols[i].col] = 0;
				}
			}
		}
		for (uint32_t i = 0; i < n_cols && !need.empty(); ++i) { // a column listed twice, once in place: its chunks are needed
			if (!cp[i].compact) {
				for (uint32_t rr = 0; rr < n; ++rr) {
					need[(size_t)rr * h.n_columns + cols[i].col] = 1;
				}
			}
		}
		{
			std::lock_guard<std::mutex> g(d->mu);
			status = stage_range(t, d, rg_begin + r, n, col_ids.data(), (uint32_t)col_ids.size(), s, need.empty() ? nullptr : need.data());
		}
		uint8_t* base = static_cast<uint8_t*>(sc->out[slab & 1]);
		if (n_filters && status == FLSGPU_OK) { // FilterExecutor::Apply for the slab: selection mask -> host
			status = flsgpu_filter_mask(t, filters, n_filters, rg_begin + r, n, nullptr, reinterpret_cast<uint32_t*>(base + dev_mask),
			                            rg_mask_bytes * n, s);
			d2h(reinterpret_cast<uint8_t*>(mask_host) + (uint64_t)r * rg_mask_bytes, base + dev_mask, rg_mask_bytes * n, s);
		}
		for (uint32_t i = 0; i < n_cols && status == FLSGPU_OK; ++i) {
			const flsgpu_scan_column& c = cols[i];
			const ColPlan& p = cp[i];
			uint8_t* dev = base + p.dev_rows;
			uint8_t* host_rows = static_cast<uint8_t*>(c.out_host) + (uint64_t)r * p.rg_bytes;
			if (p.compact) {
				const std::vector<uint32_t>& vm = vec_maps[map_of[i]];
				const uint64_t slot0 = vm.empty() ? 0 : p.forms[(size_t)r * h.vec_per_rowgroup + vm[0]].flat_slot;
				if (!vm.empty()) {
					uint32_t* map_dev = reinterpret_cast<uint32_t*>(base + p.dev_map);
					uint32_t* map_pinned = sc->map_host[slab & 1] + map_used;
					map_used += vm.size();
					std::memcpy(map_pinned, vm.data(), vm.size() * 4u);
					if (cudaMemcpyAsync(map_dev, map_pinned, vm.size() * 4u, cudaMemcpyHostToDevice, s) != cudaSuccess) {
						status = cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(vector map)");
						break;
					}
					g_h2d_bytes.fetch_add(vm.size() * 4u, std::memory_order_relaxed);
					status = decode_column_impl(t, d, c.col, rg_begin + r, n, p.os, dev, s, nullptr, nullptr, map_dev, (uint32_t)vm.size());
					d2h(static_cast<uint8_t*>(c.out_host) + slot0 * FLS_VEC_SZ * p.os.ob, dev, (uint64_t)vm.size() * FLS_VEC_SZ * p.os.ob, s);
				}
				continue;
			}
			if (!p.is_string) {
				status = decode_column_impl(t, d, c.col, rg_begin + r, n, p.os, dev, s);
				d2h(host_rows, dev, p.rg_bytes * n, s);
				continue;
			}
			const uint64_t hoff = heap_off(c.col, r), hbytes = heap_off(c.col, r + n) - hoff;
			uint8_t* host_heap = static_cast<uint8_t*>(c.heap_host) + hoff;
			uint8_t* dev_heap = base + p.dev_heap;
			if (p.dict) {
				const uint64_t koff = key_off(c.col, r), kbytes = (key_off(c.col, r + n) - koff) * 16u;
				uint8_t* dev_keys = base + p.dev_keys;
				status = flsgpu_decode_string_dict(t, c.col, rg_begin + r, n, dev_keys, kbytes + 16, dev_heap, hbytes + 16,
				                                   reinterpret_cast<uint64_t>(host_heap), s);
				if (status == FLSGPU_OK) {
					status = flsgpu_decode_string_codes(t, c.col, rg_begin + r, n, reinterpret_cast<uint32_t*>(dev), p.rg_bytes * n, s);
				}
				d2h(static_cast<uint8_t*>(c.keys_host) + koff * 16u, dev_keys, kbytes, s);
			} else {
				status = flsgpu_decode_strings(t, c.col, rg_begin + r, n, dev, p.rg_bytes * n, dev_heap, hbytes + 16,
				                               reinterpret_cast<uint64_t>(host_heap), s);
			}
			d2h(host_rows, dev, p.rg_bytes * n, s);
			d2h(host_heap, dev_heap, hbytes, s);
		}
	}
	// always drain both streams before the scratch goes back to the pool
	const cudaError_t e0 = cudaStreamSynchronize(sc->streams[0]), e1 = cudaStreamSynchronize(sc->streams[1]);
	if (status == FLSGPU_OK && (e0 != cudaSuccess || e1 != cudaSuccess)) {
		status = cuda_fail(e0 != cudaSuccess ? e0 : e1, "cudaStreamSynchronize");
	}
	return status;
}

// The filtered scan with the selection applied ON THE DEVICE: per slab the filter mask is turned into per-vector output
// offsets (fls_select_plan_kernel), the slab's row count comes back to the host (4 bytes, one stream synchronisation), then
// every projected column is decoded in HBM, packed (fls_select_compact_kernel) and only the selected rows are copied out,
// back to back over the whole range.  A Q6-shaped scan (2 % of the rows pass) moves 2 % of the bytes of the masked form.
int flsgpu_scan_selected_to_host(flsgpu_table* t, const flsgpu_scan_column* cols, uint32_t n_cols, const flsgpu_filter* filters,
                                 uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg, uint32_t* row_index_host,
                                 uint64_t row_index_capacity_bytes, uint64_t* n_selected_out) {
	const char* who = "flsgpu_scan_selected_to_host";
	if (!t || (!cols && n_cols) || !filters || n_filters == 0 || !n_selected_out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_selected_to_host: null argument or no filters");
	}
	int rc = check_range(t, 0, rg_begin, n_rg, who);
	if (!rc) {
		rc = check_filters(t, filters, n_filters, who);
	}
	if (rc) {
		return rc;
	}
	const fls_file_header& h = t->header;
	const uint64_t rg_rows = (uint64_t)h.vec_per_rowgroup * FLS_VEC_SZ;
	if (rg_rows * n_rg > 0xffffffffull) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_scan_selected_to_host: more than 2^32 rows in one call");
	}
	if (row_index_host && row_index_capacity_bytes < rg_rows * n_rg * 4u) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_selected_to_host: row index buffer too small");
	}
	struct ColPlan {
		OutSpec os;
		bool is_string = false, dict = false;
		uint32_t ob = 0; // bytes per row
	};
	std::vector<ColPlan> cp(n_cols);
	uint32_t ob_max = row_index_host ? 4u : 0u;
	for (uint32_t i = 0; i < n_cols; ++i) {
		const flsgpu_scan_column& c = cols[i];
		rc = check_range(t, c.col, rg_begin, n_rg, who);
		if (rc) {
			return rc;
		}
		ColPlan& p = cp[i];
		p.is_string = t->columns[c.col].phys_bits == 128;
		if (p.is_string) {
			p.dict = c.string_form == FLSGPU_STRING_DICTIONARY;
			uint64_t heap_need = 0, keys_need = 0;
			flsgpu_string_heap_bytes(t, c.col, rg_begin, n_rg, &heap_need);
			if (p.dict) {
				rc = flsgpu_string_dict_keys(t, c.col, rg_begin, n_rg, &keys_need); // fails unless every chunk is dictionary encoded
				if (rc) {
					return rc;
				}
				if (keys_need > 0xffffffffull) {
					return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_scan_selected_to_host: more than 2^32 dictionary keys in the range");
				}
			}
			if (!c.heap_host || c.heap_capacity_bytes < heap_need || (p.dict && (!c.keys_host || c.keys_capacity_bytes < keys_need * 16u))) {
				return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_selected_to_host: string heap / key buffer null or too small");
			}
			p.ob = p.dict ? 4 : 16;
		} else {
			rc = resolve_out(t->columns[c.col], c.out_type, &p.os);
			if (rc) {
				return rc;
			}
			p.ob = p.os.ob;
		}
		if (!c.out_host || c.out_capacity_bytes < rg_rows * n_rg * p.ob) { // how many rows pass is not known beforehand
			return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_selected_to_host: output buffer null or smaller than the unfiltered column");
		}
		ob_max = std::max(ob_max, p.ob);
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	auto heap_off = [&](uint32_t col, uint32_t r) {
		const uint64_t i0 = (uint64_t)rg_begin * h.n_columns + col;
		return d->heap_prefix[i0 + (uint64_t)r * h.n_columns] - d->heap_prefix[i0];
	};
	// slab: decode area A + packed area B (the widest column each) + the largest string heap of a slab + mask + offsets
	const uint64_t slab_target = (uint64_t)env_u32("FLSGPU_SLAB_MB", 256, 1, 65536) << 20;
	const uint32_t slab_rg = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(n_rg, slab_target / std::max<uint64_t>(1, rg_rows * std::max(1u, ob_max) * 2u)));
	auto key_off = [&](uint32_t col, uint32_t r) {
		const uint64_t i0 = (uint64_t)rg_begin * h.n_columns + col;
		return d->key_prefix[i0 + (uint64_t)r * h.n_columns] - d->key_prefix[i0];
	};
	uint64_t heap_max = 0, keys_max = 0;
	for (uint32_t i = 0; i < n_cols; ++i) {
		if (cp[i].is_string) {
			for (uint32_t r = 0; r < n_rg; r += slab_rg) {
				const uint32_t e = std::min(n_rg, r + slab_rg);
				heap_max = std::max(heap_max, heap_off(cols[i].col, e) - heap_off(cols[i].col, r));
				if (cp[i].dict) {
					keys_max = std::max(keys_max, key_off(cols[i].col, e) - key_off(cols[i].col, r));
				}
			}
		}
	}
	const uint64_t area = (rg_rows * slab_rg * std::max(1u, ob_max) + 255u) & ~255ull;
	const uint64_t dev_a = 0, dev_b = area, dev_heap = 2u * area, dev_keys = dev_heap + ((heap_max + 16u + 255u) & ~255ull);
	const uint64_t dev_mask = dev_keys + ((keys_max * 16u + 16u + 255u) & ~255ull);
	const uint64_t rg_mask_bytes = (uint64_t)h.vec_per_rowgroup * 128u;
	const uint64_t dev_offsets = dev_mask + rg_mask_bytes * slab_rg;
	const uint64_t slab_bytes = dev_offsets + ((((uint64_t)slab_rg * h.vec_per_rowgroup + 1u) * 4u + 255u) & ~255ull);
	FlsScanScratch* sc = nullptr;
	{
		std::lock_guard<std::mutex> g(d->mu);
		if (!d->scratch_free.empty()) {
			sc = d->scratch_free.back();
			d->scratch_free.pop_back();
		} else {
			sc = new FlsScanScratch();
			d->scratch_all.push_back(sc);
		}
	}
	struct Return { // hand the scratch back on every exit path
		FlsDeviceState* d;
		FlsScanScratch* sc;
		~Return() {
			std::lock_guard<std::mutex> g(d->mu);
			d->scratch_free.push_back(sc);
		}
	} ret {d, sc};
	if (sc->out_bytes < slab_bytes) {
		sc->out_bytes = 0;
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaFree(sc->out[i]));
			sc->out[i] = nullptr;
		}
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaMalloc(&sc->out[i], slab_bytes));
		}
		sc->out_bytes = slab_bytes;
	}
	for (int i = 0; i < 2; ++i) {
		if (!sc->streams[i]) {
			FLS_CUDA(cudaStreamCreateWithFlags(&sc->streams[i], cudaStreamNonBlocking));
		}
	}
	if (!sc->count_host) {
		FLS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&sc->count_host), 256, cudaHostAllocDefault));
	}
	std::vector<uint32_t> col_ids;
	for (uint32_t i = 0; i < n_cols; ++i) {
		if (std::find(col_ids.begin(), col_ids.end(), cols[i].col) == col_ids.end()) {
			col_ids.push_back(cols[i].col);
		}
	}
	for (uint32_t i = 0; i < n_filters; ++i) {
		if (std::find(col_ids.begin(), col_ids.end(), filters[i].col) == col_ids.end()) {
			col_ids.push_back(filters[i].col);
		}
	}
	int status = FLSGPU_OK;
	auto d2h = [&](void* dst, const void* src, uint64_t bytes, cudaStream_t s) {
		if (status == FLSGPU_OK && bytes && cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s) != cudaSuccess) {
			status = cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(D2H)");
		}
		g_d2h_bytes.fetch_add(bytes, std::memory_order_relaxed);
	};
	uint64_t written = 0; // selected rows of the slabs so far
	uint32_t slab = 0;
	for (uint32_t r = 0; r < n_rg && status == FLSGPU_OK; r += slab_rg, ++slab) {
		const uint32_t n = std::min(slab_rg, n_rg - r), n_vec = n * h.vec_per_rowgroup;
		cudaStream_t s = sc->streams[slab & 1];
		uint8_t* base = static_cast<uint8_t*>(sc->out[slab & 1]);
		uint32_t* mask_dev = reinterpret_cast<uint32_t*>(base + dev_mask);
		uint32_t* offsets_dev = reinterpret_cast<uint32_t*>(base + dev_offsets);
		{
			std::lock_guard<std::mutex> g(d->mu);
			status = stage_range(t, d, rg_begin + r, n, col_ids.data(), (uint32_t)col_ids.size(), s);
		}
		if (status == FLSGPU_OK) {
			status = flsgpu_filter_mask(t, filters, n_filters, rg_begin + r, n, nullptr, mask_dev, rg_mask_bytes * n, s);
		}
		if (status != FLSGPU_OK) {
			break;
		}
		cudaError_t ce = fls_select_launch_plan(mask_dev, n_vec, offsets_dev, s);
		if (ce == cudaSuccess) {
			ce = cudaMemcpyAsync(sc->count_host + (slab & 1), offsets_dev + n_vec, 4, cudaMemcpyDeviceToHost, s);
		}
		if (ce == cudaSuccess) {
			ce = cudaStreamSynchronize(s); // the slab's row count decides the size of every copy below
		}
		if (ce != cudaSuccess) {
			status = cuda_fail(ce, "selection plan");
			break;
		}
		g_launches.fetch_add(1, std::memory_order_relaxed);
		const uint64_t n_sel = sc->count_host[slab & 1];
		for (uint32_t i = 0; i < n_cols && status == FLSGPU_OK; ++i) {
			const flsgpu_scan_column& c = cols[i];
			const ColPlan& p = cp[i];
			uint8_t* dev = base + dev_a;
			if (n_sel == 0 && !(p.is_string && p.dict)) {
				continue; // nothing of this slab passes: no decode, no copy (dictionaries are still delivered whole)
			}
			if (p.is_string && p.dict) {
				// keys + dictionary bytes of the slab's row groups go to their places in the range-wide dictionary; the packed
				// codes count from the start of that dictionary
				const uint64_t hoff = heap_off(c.col, r), hbytes = heap_off(c.col, r + n) - hoff;
				const uint64_t koff = key_off(c.col, r), kbytes = (key_off(c.col, r + n) - koff) * 16u;
				uint8_t* host_heap = static_cast<uint8_t*>(c.heap_host) + hoff;
				status = flsgpu_decode_string_dict(t, c.col, rg_begin + r, n, base + dev_keys, kbytes + 16, base + dev_heap, hbytes + 16,
				                                   reinterpret_cast<uint64_t>(host_heap), s);
				if (status == FLSGPU_OK && n_sel) {
					status = flsgpu_decode_string_codes(t, c.col, rg_begin + r, n, reinterpret_cast<uint32_t*>(dev), rg_rows * n * 4u, s);
				}
				if (status != FLSGPU_OK) {
					break;
				}
				d2h(static_cast<uint8_t*>(c.keys_host) + koff * 16u, base + dev_keys, kbytes, s);
				d2h(host_heap, base + dev_heap, hbytes, s);
				if (n_sel == 0) {
					continue;
				}
				const uint64_t i0 = (uint64_t)rg_begin * h.n_columns + c.col;
				ce = fls_select_launch_compact_codes(reinterpret_cast<const uint32_t*>(dev), mask_dev, offsets_dev,
				                                     reinterpret_cast<uint32_t*>(base + dev_b), n_vec,
				                                     reinterpret_cast<const unsigned long long*>(d->d_key_prefix + i0), d->key_prefix[i0],
				                                     h.n_columns, h.vec_per_rowgroup, r, s);
				if (ce != cudaSuccess) {
					status = cuda_fail(ce, "fls_select_compact_kernel(codes)");
					break;
				}
				g_launches.fetch_add(1, std::memory_order_relaxed);
				d2h(static_cast<uint8_t*>(c.out_host) + written * 4u, base + dev_b, n_sel * 4u, s);
				continue;
			}
			if (p.is_string) {
				const uint64_t hoff = heap_off(c.col, r), hbytes = heap_off(c.col, r + n) - hoff;
				uint8_t* host_heap = static_cast<uint8_t*>(c.heap_host) + hoff;
				status = flsgpu_decode_strings(t, c.col, rg_begin + r, n, dev, rg_rows * n * 16u, base + dev_heap, hbytes + 16,
				                               reinterpret_cast<uint64_t>(host_heap), s);
				d2h(host_heap, base + dev_heap, hbytes, s);
			} else {
				status = decode_column_impl(t, d, c.col, rg_begin + r, n, p.os, dev, s);
			}
			if (status != FLSGPU_OK) {
				break;
			}
			ce = fls_select_launch_compact(dev, p.ob, mask_dev, offsets_dev, base + dev_b, n_vec, 0, s);
			if (ce != cudaSuccess) {
				status = cuda_fail(ce, "fls_select_compact_kernel");
				break;
			}
			g_launches.fetch_add(1, std::memory_order_relaxed);
			d2h(static_cast<uint8_t*>(c.out_host) + written * p.ob, base + dev_b, n_sel * p.ob, s);
		}
		if (row_index_host && n_sel && status == FLSGPU_OK) {
			ce = fls_select_launch_compact(nullptr, 4, mask_dev, offsets_dev, base + dev_b, n_vec, (uint32_t)(r * rg_rows), s);
			if (ce != cudaSuccess) {
				status = cuda_fail(ce, "fls_select_compact_kernel(row index)");
				break;
			}
			g_launches.fetch_add(1, std::memory_order_relaxed);
			d2h(row_index_host + written, base + dev_b, n_sel * 4u, s);
		}
		written += n_sel;
	}
	// always drain both streams before the scratch goes back to the pool
	const cudaError_t e0 = cudaStreamSynchronize(sc->streams[0]), e1 = cudaStreamSynchronize(sc->streams[1]);
	if (status == FLSGPU_OK && (e0 != cudaSuccess || e1 != cudaSuccess)) {
		status = cuda_fail(e0 != cudaSuccess ? e0 : e1, "cudaStreamSynchronize");
	}
	if (status == FLSGPU_OK) {
		*n_selected_out = written;
	}
	return status;
}

int flsgpu_scan_columns_to_host(flsgpu_table* t, const uint32_t* col_ids, uint32_t n_cols, uint32_t rg_begin, uint32_t n_rg,
                                const int32_t* out_types, void* const* out_hosts, const uint64_t* out_capacity_bytes) {
	if (!t || !col_ids || !out_hosts || !out_capacity_bytes || n_cols == 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_columns_to_host: null argument");
	}
	std::vector<flsgpu_scan_column> cols(n_cols);
	for (uint32_t i = 0; i < n_cols; ++i) {
		if (col_ids[i] < t->header.n_columns && t->columns[col_ids[i]].phys_bits == 128) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_scan_columns_to_host: string column (use flsgpu_scan_projection_to_host)");
		}
		cols[i] = flsgpu_scan_column {};
		cols[i].col = col_ids[i];
		cols[i].out_type = out_types ? out_types[i] : FLSGPU_OUT_NATIVE;
		cols[i].out_host = out_hosts[i];
		cols[i].out_capacity_bytes = out_capacity_bytes[i];
	}
	return flsgpu_scan_projection_to_host(t, cols.data(), n_cols, rg_begin, n_rg);
}

int flsgpu_scan_column_to_host(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, int32_t out_type, void* out_host,
                               uint64_t out_capacity_bytes) {
	return flsgpu_scan_columns_to_host(t, &col, 1, rg_begin, n_rg, &out_type, &out_host, &out_capacity_bytes);
}

int flsgpu_sync(void* cuda_stream) {
	if (cuda_stream == nullptr) { // the default stream is per device: the caller's current device, or our default
		int cur = -1;
		if (cudaGetDevice(&cur) != cudaSuccess) {
			const int rc = ensure_device();
			if (rc) {
				return rc;
			}
		}
	}
	FLS_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(cuda_stream)));
	return FLSGPU_OK;
}

// Pinned allocations are cached: page-locking costs ~0.3 ms per MiB, and every scan of every query wants staging buffers of
// the same few sizes (a worker's batch of row groups).  Freed blocks go to a free list (up to FLSGPU_HOST_POOL_MB, default
// 8192, in total) and are handed out again to requests they fit without wasting more than half of the block.
namespace {
struct HostPool {
	std::mutex mu;
	std::map<void*, uint64_t> live;                  // block -> its real size
	std::multimap<uint64_t, void*> free_blocks;      // size -> block
	uint64_t cached = 0;
};
HostPool& host_pool() {
	static HostPool* p = new HostPool(); // never destroyed: blocks may be returned during static destruction
	return *p;
}
} // namespace

int flsgpu_host_alloc(uint64_t bytes, void** out) {
	if (!out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_host_alloc: null argument");
	}
	const int rc = ensure_device();
	if (rc) {
		return rc;
	}
	const uint64_t want = bytes ? bytes : 1;
	HostPool& hp = host_pool();
	{
		std::lock_guard<std::mutex> g(hp.mu);
		auto it = hp.free_blocks.lower_bound(want);
		if (it != hp.free_blocks.end() && it->first <= std::max<uint64_t>(2 * want, 1u << 20)) {
			*out = it->second;
			hp.cached -= it->first;
			hp.free_blocks.erase(it);
			return FLSGPU_OK;
		}
	}
	FLS_CUDA(cudaHostAlloc(out, want, cudaHostAllocPortable));
	std::lock_guard<std::mutex> g(hp.mu);
	hp.live[*out] = want;
	return FLSGPU_OK;
}

int flsgpu_host_free(void* p) {
	if (!p) {
		return FLSGPU_OK;
	}
	static const uint64_t limit = (uint64_t)env_u32("FLSGPU_HOST_POOL_MB", 8192, 0, 1u << 20) << 20;
	HostPool& hp = host_pool();
	{
		std::lock_guard<std::mutex> g(hp.mu);
		auto it = hp.live.find(p);
		if (it != hp.live.end() && hp.cached + it->second <= limit) {
			hp.free_blocks.emplace(it->second, p);
			hp.cached += it->second;
			return FLSGPU_OK;
		}
		if (it != hp.live.end()) {
			hp.live.erase(it);
		}
	}
	FLS_CUDA(cudaFreeHost(p));
	return FLSGPU_OK;
}

void flsgpu_transfer_bytes(uint64_t* h2d, uint64_t* d2h) {
	if (h2d) {
		*h2d = g_h2d_bytes.load(std::memory_order_relaxed);
	}
	if (d2h) {
		*d2h = g_d2h_bytes.load(std::memory_order_relaxed);
	}
}

uint64_t flsgpu_launch_count(void) {
	return g_launches.load(std::memory_order_relaxed);
}

int flsgpu_aggregate_column(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, flsgpu_agg_result* result_dev,
                            void* cuda_stream) {
	int rc = check_range(t, col, rg_begin, n_rg, "flsgpu_aggregate_column");
	if (rc) {
		return rc;
	}
	if (!result_dev || (reinterpret_cast<uintptr_t>(result_dev) & 7u)) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_aggregate_column: result pointer null or misaligned");
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	rc = check_resident(t, d, col, rg_begin, n_rg, static_cast<cudaStream_t>(cuda_stream));
	if (rc) {
		return rc;
	}
	OutSpec os;
	os.ob = t->columns[col].phys_bits / 8;
	os.conv = 0;
	cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
	flsgpu_agg_acc* acc = reinterpret_cast<flsgpu_agg_acc*>(result_dev);
	rc = decode_column_impl(t, d, col, rg_begin, n_rg, os, nullptr, s, acc);
	if (rc) {
		return rc;
	}
	// MIN / MAX come from the chunks' zone maps (exact, written from the data: reference row_group_statistics.cpp:60-127);
	// integer columns only, as order keys (value ^ sign bit) so that signed and unsigned columns merge the same way
	unsigned long long key_min = ~0ull, key_max = 0ull;
	if (!t->columns[col].is_float) {
		const unsigned long long flip = t->columns[col].is_signed ? 0x8000000000000000ull : 0ull;
		for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
			const fls_chunk_desc& c = t->chunk(r, col);
			if (c.has_stats) {
				key_min = std::min<unsigned long long>(key_min, c.min_val ^ flip);
				key_max = std::max<unsigned long long>(key_max, c.max_val ^ flip);
			}
		}
	}
	FLS_CUDA(fls_agg_finalize(acc, t->columns[col].is_signed, key_min, key_max, s));
	return FLSGPU_OK;
}

// ---- string columns ------------------------------------------------------------------------------------------------
namespace {

int string_launch_setup(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const char* who, FlsDeviceState** dout,
                        FlsStrLaunch* L, cudaStream_t stream) {
	int rc = check_range(t, col, rg_begin, n_rg, who);
	if (rc) {
		return rc;
	}
	if (t->columns[col].phys_bits != 128) {
		char msg[160];
		std::snprintf(msg, sizeof(msg), "%s: column %u is not a string column", who, col);
		return fls_set_error(FLSGPU_ERR_INVALID, msg);
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	rc = check_resident(t, d, col, rg_begin, n_rg, stream);
	if (rc) {
		return rc;
	}
	const uint64_t i0 = (uint64_t)rg_begin * t->header.n_columns + col;
	L->chunks = d->d_chunks + i0;
	L->heap_prefix = d->d_heap_prefix + i0;
	L->heap_prefix0 = d->heap_prefix[i0];
	L->key_prefix = d->d_key_prefix + i0;
	L->key_prefix0 = d->key_prefix[i0];
	L->dict_only = 0;
	L->fuse_rows = 0;
	L->pdl = g_pdl ? 1u : 0u;
	// FSST tiling: size a warp tile (32 * lane_strings strings) so that its codes + decoded bytes take <= 8 KiB of shared
	// memory at twice the column's average string length: the decode is latency bound (dependent shared-memory lookups per
	// code), so resident warps matter more than long per-lane runs (measured, 20 M l_comment-like strings: 26 KiB/warp
	// 1.17 ms, 14 KiB 0.79 ms, 7-10 KiB 0.63 ms, <= 5 KiB (one string per lane) 0.70 ms)
	uint64_t n_str = 0, dec_bytes = 0, comp_bytes = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		const fls_chunk_desc& ch = t->chunk(r, col);
		if (FLS_ENC_IS_FSST_ANY(ch.encoding)) {
			n_str += ch.aux_count;
			dec_bytes += ch.const_val;
			comp_bytes += ch.seg_size[FLS_SEG_AUX0] - FLS_FSST_TABLE_BYTES(ch.encoding);
		}
	}
	L->lane_strings = 8;
	// experiment, off: lanes splitting a tile by code bytes instead of by strings evens out the lanes' work but costs a lower-bound
	// search per lane and tile; measured slower (dbgen l_comment SF3 0.645 -> 0.698 ms, synthetic text 0.523 -> 0.631 ms)
	L->balance = env_u32("FLSGPU_STR_BALANCE", 0, 0, 1);
	L->cbuf_bytes = L->dbuf_bytes = 0;
	uint32_t max_fsst_strings = 0;
	uint64_t max_plain_bytes = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		const fls_chunk_desc& ch = t->chunk(r, col);
		if (FLS_ENC_IS_FSST_ANY(ch.encoding)) {
			max_fsst_strings = std::max(max_fsst_strings, ch.aux_count);
		} else {
			max_plain_bytes = std::max<uint64_t>(max_plain_bytes, ch.const_val);
		}
	}
	if (n_str) {
		const double avg_dec = (double)dec_bytes / (double)n_str, avg_comp = (double)comp_bytes / (double)n_str;
		const uint32_t budget = env_u32("FLSGPU_STR_WARP_KB", 10, 2, 26) * 1024u, cap = 26u * 1024u;
		uint32_t ls = std::min(16u, env_u32("FLSGPU_STR_LANE_STRINGS", 16, 1, 16));
		for (;; ls >>= 1) {
			const uint32_t db = ((uint32_t)(2.0 * 32.0 * ls * avg_dec) + 64u + 255u) & ~255u;
			const uint32_t cb = ((uint32_t)(2.0 * 32.0 * ls * avg_comp) + 64u + 255u) & ~255u;
			if (db + 2u * cb <= budget || ls == 1) { // two staging buffers for codes (the next tile's are prefetched) + one decode buffer
				L->lane_strings = ls;
				L->dbuf_bytes = db;
				L->cbuf_bytes = cb;
				if (db + 2u * cb > cap) { // very long strings: whatever fits goes through the buffers, the rest decodes to global
					L->dbuf_bytes = (cap / 2u) & ~255u;
					L->cbuf_bytes = (cap / 4u) & ~255u;
				}
				break;
			}
		}
	}
	// CTAs per chunk of the heap kernel: enough 8-warp CTAs for the largest FSST chunk's tiles, and one CTA per 32 KiB of
	// plain bytes (a 9 KiB dictionary is one CTA's work, not 32 CTAs copying 280 bytes each)
	// (a warp of the heap kernel walks FLSGPU_STR_TILES tiles of the largest chunk as a software pipeline)
	const uint32_t per_cta = 8u * 32u * L->lane_strings * std::max(1u, env_u32("FLSGPU_STR_TILES", 8, 1, 1024));
	L->groups_per_rg = std::max<uint32_t>(1u, std::max<uint32_t>((max_fsst_strings + per_cta - 1u) / per_cta,
	                                                            (uint32_t)((max_plain_bytes + 32767u) >> 15)));
	L->chunk_stride = t->header.n_columns;
	L->vec_per_rg = t->header.vec_per_rowgroup;
	L->n_rg = n_rg;
	L->codes_only = 0;
	*dout = d;
	return FLSGPU_OK;
}

bool is_dict_string_enc(unsigned enc) {
	return FLS_ENC_IS_STR_DICT(enc);
}

} // namespace

int flsgpu_decode_strings(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_strings_dev,
                          uint64_t out_capacity_bytes, void* heap_dev, uint64_t heap_capacity_bytes, uint64_t heap_base_addr,
                          void* cuda_stream) {
	FlsDeviceState* d = nullptr;
	FlsStrLaunch L {};
	int rc = string_launch_setup(t, col, rg_begin, n_rg, "flsgpu_decode_strings", &d, &L, static_cast<cudaStream_t>(cuda_stream));
	if (rc) {
		return rc;
	}
	uint64_t heap_need = 0;
	flsgpu_string_heap_bytes(t, col, rg_begin, n_rg, &heap_need);
	const uint64_t need = (uint64_t)n_rg * t->header.vec_per_rowgroup * FLS_VEC_SZ * 16u;
	if (!out_strings_dev || (reinterpret_cast<uintptr_t>(out_strings_dev) & 15u) || out_capacity_bytes < need) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_decode_strings: string_t buffer null, misaligned or too small");
	}
	if (!heap_dev || (reinterpret_cast<uintptr_t>(heap_dev) & 15u) || heap_capacity_bytes < heap_need) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_decode_strings: heap null, misaligned or smaller than flsgpu_string_heap_bytes");
	}
	L.heap = static_cast<uint8_t*>(heap_dev);
	L.heap_base_addr = heap_base_addr;
	L.out = out_strings_dev;
	L.fuse_rows = env_u32("FLSGPU_STR_FUSE_ROWS", 1, 0, 1);
	cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
	FLS_CUDA(fls_str_launch_heap(L, s));
	FLS_CUDA(fls_str_launch_rows(L, s));
	g_launches.fetch_add(2, std::memory_order_relaxed);
	return FLSGPU_OK;
}

int flsgpu_decode_string_codes(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint32_t* out_codes_dev,
                               uint64_t out_capacity_bytes, void* cuda_stream) {
	FlsDeviceState* d = nullptr;
	FlsStrLaunch L {};
	int rc = string_launch_setup(t, col, rg_begin, n_rg, "flsgpu_decode_string_codes", &d, &L, static_cast<cudaStream_t>(cuda_stream));
	if (rc) {
		return rc;
	}
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		if (!is_dict_string_enc(t->chunk(r, col).encoding)) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_decode_string_codes: chunk is not dictionary encoded (decode it with flsgpu_decode_strings)");
		}
	}
	const uint64_t need = (uint64_t)n_rg * t->header.vec_per_rowgroup * FLS_VEC_SZ * 4u;
	if (!out_codes_dev || (reinterpret_cast<uintptr_t>(out_codes_dev) & 15u) || out_capacity_bytes < need) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_decode_string_codes: output buffer null, misaligned or too small");
	}
	L.out = out_codes_dev;
	L.codes_only = 1;
	FLS_CUDA(fls_str_launch_rows(L, static_cast<cudaStream_t>(cuda_stream)));
	g_launches.fetch_add(1, std::memory_order_relaxed);
	return FLSGPU_OK;
}

int flsgpu_decode_string_dict(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_keys_dev,
                              uint64_t keys_capacity_bytes, void* heap_dev, uint64_t heap_capacity_bytes, uint64_t heap_base_addr,
                              void* cuda_stream) {
	FlsDeviceState* d = nullptr;
	FlsStrLaunch L {};
	int rc = string_launch_setup(t, col, rg_begin, n_rg, "flsgpu_decode_string_dict", &d, &L, static_cast<cudaStream_t>(cuda_stream));
	if (rc) {
		return rc;
	}
	uint64_t n_keys = 0, heap_need = 0;
	uint32_t max_keys = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		const fls_chunk_desc& ch = t->chunk(r, col);
		if (!is_dict_string_enc(ch.encoding)) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_decode_string_dict: chunk is not dictionary encoded (decode it with flsgpu_decode_strings)");
		}
		n_keys += ch.aux_count;
		max_keys = std::max(max_keys, ch.aux_count);
		heap_need += (ch.const_val + 15u) & ~15ull;
	}
	if (!out_keys_dev || (reinterpret_cast<uintptr_t>(out_keys_dev) & 15u) || keys_capacity_bytes < n_keys * 16u) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_decode_string_dict: key buffer null, misaligned or too small");
	}
	if (!heap_dev || (reinterpret_cast<uintptr_t>(heap_dev) & 15u) || heap_capacity_bytes < heap_need) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_decode_string_dict: heap null, misaligned or smaller than flsgpu_string_heap_bytes");
	}
	L.heap = static_cast<uint8_t*>(heap_dev);
	L.heap_base_addr = heap_base_addr;
	L.out = out_keys_dev;
	L.dict_only = 1;
	// plain dictionaries (DICT_STR): the heap launch builds the keys' string_t as well - one launch instead of two (the keys
	// launch had to wait for the whole heap; FSST-compressed dictionaries still need it: their keys exist only decoded)
	bool plain = env_u32("FLSGPU_STR_FUSE_KEYS", 1, 0, 1) != 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg && plain; ++r) {
		plain = t->chunk(r, col).encoding == FLS_ENC_DICT_STR;
	}
	L.fuse_keys = plain ? 1u : 0u;
	cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
	FLS_CUDA(fls_str_launch_heap(L, s));
	if (!plain) {
		FLS_CUDA(fls_str_launch_keys(L, max_keys, s));
	}
	g_launches.fetch_add(plain ? 1 : 2, std::memory_order_relaxed);
	return FLSGPU_OK;
}

int flsgpu_scan_strings_to_host(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_strings_host,
                                uint64_t out_capacity_bytes, void* heap_host, uint64_t heap_capacity_bytes) {
	if (!t || col >= t->header.n_columns || t->columns[col].phys_bits != 128) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_scan_strings_to_host: not a string column");
	}
	flsgpu_scan_column c {};
	c.col = col;
	c.string_form = FLSGPU_STRING_FLAT;
	c.out_host = out_strings_host;
	c.out_capacity_bytes = out_capacity_bytes;
	c.heap_host = heap_host;
	c.heap_capacity_bytes = heap_capacity_bytes;
	return flsgpu_scan_projection_to_host(t, &c, 1, rg_begin, n_rg);
}

// ---- filter pushdown ---------------------------------------------------------------------------------------------------
namespace {

// order of two physical values of a column (integers in their signedness; floats with DuckDB's NaN rules).  Zone maps of
// float columns are NOT physical words (they are doubles, fls_format.h): pruning skips float columns and never comes here with one.
int cmp_physical(const fls_column_desc& c, uint64_t a, uint64_t b) {
	if (c.is_float) {
		double x, y;
		if (c.phys_bits == 32) {
			float fx, fy;
			const uint32_t ua = (uint32_t)a, ub = (uint32_t)b;
			std::memcpy(&fx, &ua, 4);
			std::memcpy(&fy, &ub, 4);
			x = fx;
			y = fy;
		} else {
			std::memcpy(&x, &a, 8);
			std::memcpy(&y, &b, 8);
		}
		const bool xn = x != x, yn = y != y;
		if (xn || yn) {
			return xn && yn ? 0 : xn ? 1 : -1;
		}
		return x < y ? -1 : x > y ? 1 : 0;
	}
	if (c.is_signed) {
		return (int64_t)a < (int64_t)b ? -1 : (int64_t)a > (int64_t)b ? 1 : 0;
	}
	return a < b ? -1 : a > b ? 1 : 0;
}

// Stream-ordered scratch of the string filters: a pool of its own (the device's default pool hands memory back to the
// driver at every synchronisation, which made each filter pass pay for fresh pages: 0.46 ms per 512 MiB slab); it keeps up
// to FLSGPU_SCRATCH_POOL_MB (default 2048) of freed blocks for the next pass.
cudaError_t scratch_alloc(void** out, uint64_t bytes, cudaStream_t s) {
	static std::mutex mu;
	static cudaMemPool_t pools[64] = {};
	int dev = 0;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) {
		return e;
	}
	if (dev < 0 || dev >= 64) {
		return cudaMallocAsync(out, bytes, s);
	}
	{
		std::lock_guard<std::mutex> g(mu);
		if (!pools[dev]) {
			cudaMemPoolProps props {};
			props.allocType = cudaMemAllocationTypePinned;
			props.handleTypes = cudaMemHandleTypeNone;
			props.location.type = cudaMemLocationTypeDevice;
			props.location.id = dev;
			e = cudaMemPoolCreate(&pools[dev], &props);
			if (e != cudaSuccess) {
				return e;
			}
			uint64_t keep = (uint64_t)env_u32("FLSGPU_SCRATCH_POOL_MB", 2048, 0, 1u << 20) << 20;
			cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
		}
	}
	return cudaMallocFromPoolAsync(out, bytes, pools[dev], s);
}

// Dictionary-encoded ranges (DICT_STR / FSST_DICT / FSST12_DICT: l_shipmode, l_returnflag ...): the comparison is evaluated
// once per dictionary key (keys + dictionary bytes decoded into scratch, one flag byte per key), then the rows kernel unpacks
// the codes and looks the flags up: per row only the packed code is read and one mask bit written.
int string_filter_pass_dict(flsgpu_table* t, const flsgpu_filter& f, uint32_t rg_begin, uint32_t n_rg, const uint8_t* rg_keep_dev,
                            uint32_t* mask_dev, bool and_into, cudaStream_t s) {
	uint64_t n_keys = 0, heap_bytes = 0;
	int rc = flsgpu_string_dict_keys(t, f.col, rg_begin, n_rg, &n_keys);
	if (!rc) {
		rc = flsgpu_string_heap_bytes(t, f.col, rg_begin, n_rg, &heap_bytes);
	}
	if (rc) {
		return rc;
	}
	if (n_keys > 0xffffffffull) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "string filter: more than 2^32 dictionary keys in the range");
	}
	const uint64_t keys_cap = ((n_keys * 16u + 255u) & ~255ull) + 256u, heap_cap = ((heap_bytes + 255u) & ~255ull) + 256u;
	const uint64_t flags_cap = ((n_keys + 255u) & ~255ull) + 256u, const_cap = ((f.str_len + 255u) & ~255ull) + 256u;
	uint8_t* scratch = nullptr;
	FLS_CUDA(scratch_alloc(reinterpret_cast<void**>(&scratch), keys_cap + heap_cap + flags_cap + const_cap + (uint64_t)n_rg * 4u, s));
	uint8_t* keys_dev = scratch;
	uint8_t* heap_dev = keys_dev + keys_cap;
	uint8_t* flags_dev = heap_dev + heap_cap;
	uint8_t* const_dev = flags_dev + flags_cap;
	uint32_t* counts_dev = reinterpret_cast<uint32_t*>(const_dev + const_cap);
	std::vector<uint32_t> counts(n_rg);
	for (uint32_t r = 0; r < n_rg; ++r) {
		counts[r] = t->rowgroups[rg_begin + r].n_tuples;
	}
	cudaError_t ce = cudaMemcpyAsync(counts_dev, counts.data(), (size_t)n_rg * 4u, cudaMemcpyHostToDevice, s); // pageable: staged
	if (ce == cudaSuccess && f.str_len) {                                                                      // before the return
		ce = cudaMemcpyAsync(const_dev, f.str_data, f.str_len, cudaMemcpyHostToDevice, s);
	}
	if (ce != cudaSuccess) {
		rc = fls_set_error(FLSGPU_ERR_CUDA, cudaGetErrorString(ce));
	}
	if (!rc) {
		rc = flsgpu_decode_string_dict(t, f.col, rg_begin, n_rg, keys_dev, keys_cap, heap_dev, heap_cap, reinterpret_cast<uint64_t>(heap_dev), s);
	}
	if (!rc) {
		FlsStrFilterLaunch K {};
		K.rows = keys_dev;
		K.const_data = const_dev;
		K.const_len = (uint32_t)f.str_len;
		if (f.str_len) {
			std::memcpy(&K.const_prefix, f.str_data, (size_t)std::min<uint64_t>(4, f.str_len));
		}
		K.op = (uint32_t)f.op;
		K.key_flags_out = flags_dev;
		K.n_keys = (uint32_t)n_keys;
		ce = fls_str_launch_key_flags(K, s);
		if (ce != cudaSuccess) {
			rc = fls_set_error(FLSGPU_ERR_CUDA, cudaGetErrorString(ce));
		}
	}
	if (!rc) {
		FlsDeviceState* d = nullptr;
		FlsStrLaunch L {};
		rc = string_launch_setup(t, f.col, rg_begin, n_rg, "flsgpu_filter_mask", &d, &L, s);
		if (!rc) {
			L.key_flags = flags_dev;
			L.n_tuples = counts_dev;
			L.rg_keep = rg_keep_dev;
			L.filter_mask = mask_dev;
			L.filter_and = and_into ? 1u : 0u;
			ce = fls_str_launch_rows(L, s);
			if (ce != cudaSuccess) {
				rc = fls_set_error(FLSGPU_ERR_CUDA, cudaGetErrorString(ce));
			}
			g_launches.fetch_add(2, std::memory_order_relaxed);
		}
	}
	cudaFreeAsync(scratch, s);
	return rc;
}

// FilterExecutor::Apply on a VARCHAR / BLOB column: the reference decodes the chunk and hands the vector to
// ColumnSegment::FilterSelection (filter_executor.cpp:42-48).  Here slabs of row groups are decoded to string_t rows + heap
// in stream-ordered scratch memory (string pointers = device addresses of the heap) and compared by fls_str_filter_kernel;
// the scratch (rows, heap, the constant's bytes, the slab's tuple counts) is one cudaMallocAsync block reused by every slab.
int string_filter_pass(flsgpu_table* t, const flsgpu_filter& f, uint32_t rg_begin, uint32_t n_rg, const uint8_t* rg_keep_dev,
                       uint32_t* mask_dev, bool and_into, cudaStream_t s) {
	const uint32_t vpr = t->header.vec_per_rowgroup;
	bool all_dict = n_rg > 0 && env_u32("FLSGPU_STR_FILTER_DICT", 1, 0, 1) != 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg && all_dict; ++r) {
		all_dict = is_dict_string_enc(t->chunk(r, f.col).encoding);
	}
	if (all_dict) {
		return string_filter_pass_dict(t, f, rg_begin, n_rg, rg_keep_dev, mask_dev, and_into, s);
	}
	const uint64_t rows_per_rg = (uint64_t)vpr * FLS_VEC_SZ * 16u;
	const uint64_t budget = (uint64_t)env_u32("FLSGPU_STR_FILTER_SLAB_MB", 512, 1, 65536) << 20;
	// slabs: as many row groups as fit the budget (rows + heap), at least one
	std::vector<std::pair<uint32_t, uint32_t>> slabs; // (first row group, count)
	std::vector<uint64_t> slab_heap;
	uint64_t max_rows = 0, max_heap = 0;
	for (uint32_t r = rg_begin; r < rg_begin + n_rg;) {
		uint64_t bytes = 0, heap = 0;
		uint32_t n = 0;
		while (r + n < rg_begin + n_rg) {
			uint64_t h = 0;
			const int hrc = flsgpu_string_heap_bytes(t, f.col, r + n, 1, &h);
			if (hrc) {
				return hrc;
			}
			if (n && bytes + rows_per_rg + h > budget) {
				break;
			}
			bytes += rows_per_rg + h;
			heap += h;
			++n;
		}
		slabs.emplace_back(r, n);
		slab_heap.push_back(heap);
		max_rows = std::max(max_rows, (uint64_t)n * rows_per_rg);
		max_heap = std::max(max_heap, heap);
		r += n;
	}
	if (slabs.empty()) {
		return FLSGPU_OK;
	}
	const uint64_t heap_cap = ((max_heap + 255u) & ~255ull) + 256u;
	const uint64_t const_cap = ((f.str_len + 255u) & ~255ull) + 256u;
	const uint64_t counts_cap = (((uint64_t)n_rg * 4u + 255u) & ~255ull);
	uint8_t* scratch = nullptr;
	FLS_CUDA(scratch_alloc(reinterpret_cast<void**>(&scratch), max_rows + heap_cap + const_cap + counts_cap, s));
	uint8_t* rows_dev = scratch;
	uint8_t* heap_dev = rows_dev + max_rows;
	uint8_t* const_dev = heap_dev + heap_cap;
	uint32_t* counts_dev = reinterpret_cast<uint32_t*>(const_dev + const_cap);
	std::vector<uint32_t> counts(n_rg);
	for (uint32_t r = 0; r < n_rg; ++r) {
		counts[r] = t->rowgroups[rg_begin + r].n_tuples;
	}
	int rc = FLSGPU_OK;
	cudaError_t ce = cudaMemcpyAsync(counts_dev, counts.data(), (size_t)n_rg * 4u, cudaMemcpyHostToDevice, s); // pageable source:
	if (ce == cudaSuccess && f.str_len) { // staged before the call returns
		ce = cudaMemcpyAsync(const_dev, f.str_data, f.str_len, cudaMemcpyHostToDevice, s);
	}
	if (ce != cudaSuccess) {
		rc = fls_set_error(FLSGPU_ERR_CUDA, cudaGetErrorString(ce));
	}
	uint32_t prefix = 0;
	if (f.str_len) {
		std::memcpy(&prefix, f.str_data, (size_t)std::min<uint64_t>(4, f.str_len));
	}
	for (size_t i = 0; i < slabs.size() && rc == FLSGPU_OK; ++i) {
		const uint32_t r0 = slabs[i].first, n = slabs[i].second;
		rc = flsgpu_decode_strings(t, f.col, r0, n, rows_dev, (uint64_t)n * rows_per_rg, heap_dev, heap_cap,
		                           reinterpret_cast<uint64_t>(heap_dev), s);
		if (rc) {
			break;
		}
		FlsStrFilterLaunch L {};
		L.rows = rows_dev;
		L.n_tuples = counts_dev + (r0 - rg_begin);
		L.rg_keep = rg_keep_dev ? rg_keep_dev + (r0 - rg_begin) : nullptr;
		L.const_data = const_dev;
		L.const_len = (uint32_t)f.str_len;
		L.const_prefix = prefix;
		L.op = (uint32_t)f.op;
		L.and_into = and_into ? 1u : 0u;
		L.mask = mask_dev + (uint64_t)(r0 - rg_begin) * vpr * 32u;
		L.n_rg = n;
		L.vec_per_rg = vpr;
		ce = fls_str_launch_filter(L, s);
		if (ce != cudaSuccess) {
			rc = fls_set_error(FLSGPU_ERR_CUDA, cudaGetErrorString(ce));
		}
		g_launches.fetch_add(1, std::memory_order_relaxed);
	}
	cudaFreeAsync(scratch, s);
	return rc;
}

} // namespace

int flsgpu_rowgroup_prune(const flsgpu_table* t, const flsgpu_filter* filters, uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg,
                          uint8_t* keep_out) {
	int rc = check_range(t, 0, rg_begin, n_rg, "flsgpu_rowgroup_prune");
	if (!rc) {
		rc = check_filters(t, filters, n_filters, "flsgpu_rowgroup_prune");
	}
	if (rc) {
		return rc;
	}
	if (!keep_out) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_rowgroup_prune: null output");
	}
	for (uint32_t r = 0; r < n_rg; ++r) {
		bool keep = true;
		for (uint32_t i = 0; i < n_filters && keep; ++i) {
			const fls_chunk_desc& ch = t->chunk(rg_begin + r, filters[i].col);
			const fls_column_desc& col = t->columns[filters[i].col];
			if (!ch.has_stats || col.is_float || col.phys_bits == 128) {
				continue; // no zone map: the row group may satisfy the filter (row_group_filter.cpp:157-160)
			}
			const int cmin = cmp_physical(col, ch.min_val, filters[i].constant), cmax = cmp_physical(col, ch.max_val, filters[i].constant);
			switch (filters[i].op) { // row_group_filter.cpp:177-196 (+ the EQ case, which cannot change the result)
			case FLSGPU_CMP_GT: keep = cmax > 0; break;
			case FLSGPU_CMP_GE: keep = cmax >= 0; break;
			case FLSGPU_CMP_LT: keep = cmin < 0; break;
			case FLSGPU_CMP_LE: keep = cmin <= 0; break;
			case FLSGPU_CMP_EQ: keep = cmin <= 0 && cmax >= 0; break;
			default: break;
			}
		}
		keep_out[r] = keep ? 1 : 0;
	}
	return FLSGPU_OK;
}

int flsgpu_filter_mask(flsgpu_table* t, const flsgpu_filter* filters, uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg,
                       const uint8_t* rg_keep_dev, uint32_t* mask_dev, uint64_t mask_capacity_bytes, void* cuda_stream) {
	int rc = check_range(t, 0, rg_begin, n_rg, "flsgpu_filter_mask");
	if (!rc) {
		rc = check_filters(t, filters, n_filters, "flsgpu_filter_mask");
	}
	if (rc) {
		return rc;
	}
	const uint64_t need = (uint64_t)n_rg * t->header.vec_per_rowgroup * 128u;
	if (!mask_dev || (reinterpret_cast<uintptr_t>(mask_dev) & 15u) || mask_capacity_bytes < need || n_filters == 0) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_filter_mask: mask buffer null, misaligned or too small, or no filters");
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
	// vectors nobody writes (pruned row groups, the missing vectors of a short last row group) select nothing
	FLS_CUDA(cudaMemsetAsync(mask_dev, 0, need, s));
	// one decode pass per (column, pair of comparisons): the filters are ANDed, so their order is free; a range
	// (l_shipdate >= a AND l_shipdate < b) costs one pass instead of two
	std::vector<uint8_t> done(n_filters, 0);
	bool first = true;
	for (uint32_t i = 0; i < n_filters; ++i) {
		if (done[i] || t->columns[filters[i].col].phys_bits == 128) {
			continue; // string filters run after the fused numeric passes
		}
		rc = check_resident(t, d, filters[i].col, rg_begin, n_rg, s);
		if (rc) {
			return rc;
		}
		MaskSpec ms;
		ms.mode = first ? 1u : 2u;
		first = false;
		ms.op = (uint32_t)filters[i].op;
		ms.constant = filters[i].constant;
		ms.mask = mask_dev;
		ms.rg_keep = rg_keep_dev;
		done[i] = 1;
		for (uint32_t j = i + 1; j < n_filters; ++j) {
			if (!done[j] && filters[j].col == filters[i].col) {
				ms.op2 = (uint32_t)filters[j].op;
				ms.constant2 = filters[j].constant;
				done[j] = 1;
				break;
			}
		}
		OutSpec os;
		os.ob = t->columns[filters[i].col].phys_bits / 8;
		rc = decode_column_impl(t, d, filters[i].col, rg_begin, n_rg, os, nullptr, s, nullptr, &ms);
		if (rc) {
			return rc;
		}
	}
	for (uint32_t i = 0; i < n_filters; ++i) {
		if (t->columns[filters[i].col].phys_bits == 128) {
			rc = string_filter_pass(t, filters[i], rg_begin, n_rg, rg_keep_dev, mask_dev, !first, s);
			if (rc) {
				return rc;
			}
			first = false;
		}
	}
	return FLSGPU_OK;
}

int flsgpu_aggregate_column_masked(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const uint32_t* mask_dev,
                                   const uint8_t* rg_keep_dev, flsgpu_agg_result* result_dev, void* cuda_stream) {
	int rc = check_range(t, col, rg_begin, n_rg, "flsgpu_aggregate_column_masked");
	if (rc) {
		return rc;
	}
	if (!result_dev || (reinterpret_cast<uintptr_t>(result_dev) & 7u) || !mask_dev) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_aggregate_column_masked: result or mask pointer null or misaligned");
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	rc = check_resident(t, d, col, rg_begin, n_rg, static_cast<cudaStream_t>(cuda_stream));
	if (rc) {
		return rc;
	}
	MaskSpec ms;
	ms.mode = 3;
	ms.mask = const_cast<uint32_t*>(mask_dev);
	ms.rg_keep = rg_keep_dev;
	OutSpec os;
	os.ob = t->columns[col].phys_bits / 8;
	flsgpu_agg_acc* acc = reinterpret_cast<flsgpu_agg_acc*>(result_dev);
	// SUM / COUNT only: MIN / MAX of a filtered scan cannot come from the zone maps
	return decode_column_impl(t, d, col, rg_begin, n_rg, os, nullptr, static_cast<cudaStream_t>(cuda_stream), acc, &ms);
}

// SELECT SUM(c), COUNT(*), MIN(c), MAX(c) [... WHERE <filters>] from host memory to host memory: the row groups of the range
// are worked through in slabs on two streams - H2D of the compressed chunks of the columns the query touches, (filter mask,)
// fused decode + aggregate launches - so that the copy of slab k+1 overlaps the kernels of slab k; each stream folds into a
// result set of its own, the two are added on the host.  Nothing decoded ever leaves the SMs; 64 bytes per column come back.
int flsgpu_aggregate_to_host(flsgpu_table* t, const uint32_t* col_ids, uint32_t n_cols, const flsgpu_filter* filters, uint32_t n_filters,
                             uint32_t rg_begin, uint32_t n_rg, flsgpu_agg_result* results_host) {
	if (!t || !col_ids || n_cols == 0 || !results_host) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_aggregate_to_host: null argument");
	}
	for (uint32_t i = 0; i < n_cols; ++i) {
		const int rc = check_range(t, col_ids[i], rg_begin, n_rg, "flsgpu_aggregate_to_host");
		if (rc) {
			return rc;
		}
		if (t->columns[col_ids[i]].phys_bits == 128) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_aggregate_to_host: aggregates over string columns are not supported");
		}
	}
	int rc = check_filters(t, filters, n_filters, "flsgpu_aggregate_to_host");
	if (rc) {
		return rc;
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d);
	if (rc) {
		return rc;
	}
	const fls_file_header& h = t->header;
	std::vector<uint32_t> need_cols(col_ids, col_ids + n_cols);
	for (uint32_t i = 0; i < n_filters; ++i) {
		if (std::find(need_cols.begin(), need_cols.end(), filters[i].col) == need_cols.end()) {
			need_cols.push_back(filters[i].col);
		}
	}
	std::vector<uint8_t> keep(n_rg, 1);
	if (n_filters) {
		rc = flsgpu_rowgroup_prune(t, filters, n_filters, rg_begin, n_rg, keep.data());
		if (rc) {
			return rc;
		}
	}
	const uint32_t slab_rg = std::max(1u, std::min(n_rg, env_u32("FLSGPU_AGG_SLAB_RG", 64, 1, 65535)));
	const uint64_t res_bytes = ((uint64_t)n_cols * sizeof(flsgpu_agg_result) + 255u) & ~255ull;
	const uint64_t mask_bytes = n_filters ? (uint64_t)slab_rg * h.vec_per_rowgroup * 128u : 0u;
	const uint64_t keep_bytes = n_filters ? ((uint64_t)slab_rg + 255u) & ~255ull : 0u;
	const uint64_t need_bytes = res_bytes + mask_bytes + keep_bytes;
	FlsScanScratch* sc = nullptr;
	{
		std::lock_guard<std::mutex> g(d->mu);
		if (!d->scratch_free.empty()) {
			sc = d->scratch_free.back();
			d->scratch_free.pop_back();
		} else {
			sc = new FlsScanScratch();
			d->scratch_all.push_back(sc);
		}
	}
	struct Return { // hand the scratch back on every exit path
		FlsDeviceState* d;
		FlsScanScratch* sc;
		~Return() {
			std::lock_guard<std::mutex> g(d->mu);
			d->scratch_free.push_back(sc);
		}
	} ret {d, sc};
	if (sc->out_bytes < need_bytes) {
		sc->out_bytes = 0;
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaFree(sc->out[i]));
			sc->out[i] = nullptr;
		}
		for (int i = 0; i < 2; ++i) {
			FLS_CUDA(cudaMalloc(&sc->out[i], need_bytes));
		}
		sc->out_bytes = need_bytes;
	}
	for (int i = 0; i < 2; ++i) {
		if (!sc->streams[i]) {
			FLS_CUDA(cudaStreamCreateWithFlags(&sc->streams[i], cudaStreamNonBlocking));
		}
	}
	int status = FLSGPU_OK;
	for (int i = 0; i < 2 && status == FLSGPU_OK; ++i) {
		for (uint32_t c = 0; c < n_cols && status == FLSGPU_OK; ++c) {
			status = flsgpu_agg_reset(static_cast<flsgpu_agg_result*>(sc->out[i]) + c, sc->streams[i]);
		}
	}
	uint32_t slab = 0;
	for (uint32_t r0 = 0; r0 < n_rg && status == FLSGPU_OK; r0 += slab_rg, ++slab) {
		const uint32_t n = std::min(slab_rg, n_rg - r0);
		const int b = (int)(slab & 1u);
		cudaStream_t st = sc->streams[b];
		uint8_t* base = static_cast<uint8_t*>(sc->out[b]);
		flsgpu_agg_result* res = reinterpret_cast<flsgpu_agg_result*>(base);
		{
			std::lock_guard<std::mutex> g(d->mu);
			status = stage_range(t, d, rg_begin + r0, n, need_cols.data(), (uint32_t)need_cols.size(), st);
		}
		if (status != FLSGPU_OK) {
			break;
		}
		if (n_filters) {
			uint32_t* mask_dev = reinterpret_cast<uint32_t*>(base + res_bytes);
			uint8_t* keep_dev = base + res_bytes + mask_bytes;
			if (cudaMemcpyAsync(keep_dev, keep.data() + r0, n, cudaMemcpyHostToDevice, st) != cudaSuccess) { // pageable: staged before the call returns
				status = cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(row-group keep flags)");
				break;
			}
			status = flsgpu_filter_mask(t, filters, n_filters, rg_begin + r0, n, keep_dev, mask_dev, mask_bytes, st);
			for (uint32_t c = 0; c < n_cols && status == FLSGPU_OK; ++c) {
				status = flsgpu_aggregate_column_masked(t, col_ids[c], rg_begin + r0, n, mask_dev, keep_dev, res + c, st);
			}
		} else {
			for (uint32_t c = 0; c < n_cols && status == FLSGPU_OK; ++c) {
				status = flsgpu_aggregate_column(t, col_ids[c], rg_begin + r0, n, res + c, st);
			}
		}
	}
	std::vector<flsgpu_agg_result> part(2u * n_cols);
	for (int i = 0; i < 2 && status == FLSGPU_OK; ++i) {
		if (cudaMemcpyAsync(part.data() + (size_t)i * n_cols, sc->out[i], (size_t)n_cols * sizeof(flsgpu_agg_result), cudaMemcpyDeviceToHost,
		                    sc->streams[i]) != cudaSuccess) {
			status = cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(D2H results)");
		}
	}
	// always drain both streams before the scratch goes back to the pool
	for (int i = 0; i < 2; ++i) {
		const cudaError_t e = cudaStreamSynchronize(sc->streams[i]);
		if (e != cudaSuccess && status == FLSGPU_OK) {
			status = cuda_fail(e, "cudaStreamSynchronize");
		}
	}
	if (status != FLSGPU_OK) {
		return status;
	}
	g_d2h_bytes.fetch_add(2ull * n_cols * sizeof(flsgpu_agg_result), std::memory_order_relaxed);
	for (uint32_t c = 0; c < n_cols; ++c) {
		const flsgpu_agg_result &a = part[c], &b = part[n_cols + c];
		flsgpu_agg_result& o = results_host[c];
		o = flsgpu_agg_result {};
		o.sum_lo = a.sum_lo + b.sum_lo;
		o.sum_hi = a.sum_hi + b.sum_hi + (o.sum_lo < a.sum_lo ? 1u : 0u);
		o.count = a.count + b.count;
		o.fsum = a.fsum + b.fsum;
		// order keys of the zone-map extremes (value ^ sign bit; ~0 / 0 when a stream saw none)
		const uint64_t kmin = std::min(a.reserved[0], b.reserved[0]), kmax = std::max(a.reserved[1], b.reserved[1]);
		o.reserved[0] = kmin;
		o.reserved[1] = kmax;
		if (!n_filters && o.count && kmin <= kmax && !t->columns[col_ids[c]].is_float) {
			const uint64_t flip = t->columns[col_ids[c]].is_signed ? 0x8000000000000000ull : 0ull;
			o.min_val = (int64_t)(kmin ^ flip);
			o.max_val = (int64_t)(kmax ^ flip);
		}
	}
	return FLSGPU_OK;
}

int flsgpu_agg_reset(flsgpu_agg_result* result_dev, void* cuda_stream) {
	if (!result_dev) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_agg_reset: null argument");
	}
	const int rc = ensure_device_of(result_dev);
	if (rc) {
		return rc;
	}
	FLS_CUDA(fls_agg_reset(reinterpret_cast<flsgpu_agg_acc*>(result_dev), static_cast<cudaStream_t>(cuda_stream)));
	return FLSGPU_OK;
}

} // extern "C"

// ---- allocation-free variant of the dictionary string filter ----------------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it):
// the scratch string_filter_pass_dict takes from the stream-ordered pool lives here in a file-scope device array instead.
// No counterpart in the reference.  Every device that loads the library holds
// the array until the process exits.  The array's name is a host-side stand-in in host code: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current (static_scratch_of).
#define FLS_STATIC_SCRATCH_BYTES FLSGPU_STATIC_SCRATCH_BYTES
static __device__ __align__(256) uint8_t g_static_scratch[FLS_STATIC_SCRATCH_BYTES];

// Device code names the array directly: the flag bytes of the keys start at zero on every call (the array keeps the last
// call's contents).
static __global__ void fls_static_scratch_clear_kernel(uint64_t offset, uint64_t bytes) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; // 16 bytes per thread; offset and bytes are multiples of 256
	if (i * 16u < bytes) {
		*reinterpret_cast<uint4*>(g_static_scratch + offset + i * 16u) = make_uint4(0u, 0u, 0u, 0u);
	}
}

namespace {

constexpr int kMaxStaticDevices = 64;
constexpr uint32_t kStaticMaxRowgroups = 1024; // tuple counts of the range go through a stack array

// Device address of g_static_scratch on CUDA ordinal `ordinal`, which must be the calling thread's current device.
int static_scratch_of(int ordinal, uint8_t** out) {
	static std::mutex mu;
	static uint8_t* addr[kMaxStaticDevices] = {};
	if (ordinal < 0 || ordinal >= kMaxStaticDevices) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "static scratch: device ordinal beyond 63");
	}
	std::lock_guard<std::mutex> g(mu);
	if (!addr[ordinal]) {
		int cur = -1;
		FLS_CUDA(cudaGetDevice(&cur));
		if (cur != ordinal) {
			return fls_set_error(FLSGPU_ERR_CUDA, "static scratch: the table's device is not current");
		}
		uint8_t* p = nullptr;
		FLS_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch));
		if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) {
			return fls_set_error(FLSGPU_ERR_CUDA, "static scratch: cudaGetSymbolAddress returned a null or misaligned address");
		}
		addr[ordinal] = p;
	}
	*out = addr[ordinal];
	return FLSGPU_OK;
}

} // namespace

extern "C" {

int flsgpu_filter_mask_string_dict_static(flsgpu_table* t, const flsgpu_filter* filter, uint32_t rg_begin, uint32_t n_rg,
                                          const uint8_t* rg_keep_dev, uint32_t* mask_dev, uint64_t mask_capacity_bytes,
                                          void* cuda_stream) {
	const char* who = "flsgpu_filter_mask_string_dict_static";
	int rc = check_range(t, 0, rg_begin, n_rg, who);
	if (!rc) {
		rc = check_filters(t, filter, filter ? 1u : 0u, who);
	}
	if (rc) {
		return rc;
	}
	if (!filter || t->columns[filter->col].phys_bits != 128) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_filter_mask_string_dict_static: one filter on a VARCHAR / BLOB column is required");
	}
	if (!t->dev) { // creating the device state allocates the arena: that is flsgpu_table_stage's business, not this call's
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_filter_mask_string_dict_static: table not staged (this call allocates nothing: call flsgpu_table_stage first)");
	}
	if (n_rg > kStaticMaxRowgroups) {
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_filter_mask_string_dict_static: more than 1024 row groups in one call");
	}
	const uint64_t need = (uint64_t)n_rg * t->header.vec_per_rowgroup * 128u;
	if (!mask_dev || (reinterpret_cast<uintptr_t>(mask_dev) & 15u) || mask_capacity_bytes < need) {
		return fls_set_error(FLSGPU_ERR_INVALID, "flsgpu_filter_mask_string_dict_static: mask buffer null, misaligned or too small");
	}
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		if (!is_dict_string_enc(t->chunk(r, filter->col).encoding)) {
			return fls_set_error(FLSGPU_ERR_UNSUPPORTED, "flsgpu_filter_mask_string_dict_static: a chunk of the range is not dictionary encoded (use flsgpu_filter_mask)");
		}
	}
	uint64_t n_keys = 0, heap_bytes = 0;
	rc = flsgpu_string_dict_keys(t, filter->col, rg_begin, n_rg, &n_keys);
	if (!rc) {
		rc = flsgpu_string_heap_bytes(t, filter->col, rg_begin, n_rg, &heap_bytes);
	}
	if (rc) {
		return rc;
	}
	// layout inside the array, every part 256-byte aligned: keys (string_t) | dictionary bytes | flag per key | constant | tuple counts
	const uint64_t keys_cap = ((n_keys * 16u + 255u) & ~255ull) + 256u, heap_cap = ((heap_bytes + 255u) & ~255ull) + 256u;
	const uint64_t flags_cap = ((n_keys + 255u) & ~255ull) + 256u, const_cap = ((filter->str_len + 255u) & ~255ull) + 256u;
	const uint64_t counts_cap = ((uint64_t)n_rg * 4u + 255u) & ~255ull;
	const uint64_t total = keys_cap + heap_cap + flags_cap + const_cap + counts_cap;
	if (n_keys > 0xffffffffull || total > (uint64_t)FLS_STATIC_SCRATCH_BYTES) {
		char msg[256];
		std::snprintf(msg, sizeof(msg),
		              "flsgpu_filter_mask_string_dict_static: the range needs %llu bytes of scratch (%llu keys, %llu dictionary bytes), the "
		              "static array holds %llu: use fewer row groups per call or flsgpu_filter_mask",
		              (unsigned long long)total, (unsigned long long)n_keys, (unsigned long long)heap_bytes,
		              (unsigned long long)FLS_STATIC_SCRATCH_BYTES);
		return fls_set_error(FLSGPU_ERR_UNSUPPORTED, msg);
	}
	FlsDeviceState* d = nullptr;
	rc = device_state(t, &d); // makes the table's device current; t->dev exists, nothing is created
	if (rc) {
		return rc;
	}
	uint8_t* scratch = nullptr;
	rc = static_scratch_of(d->device, &scratch);
	if (rc) {
		return rc;
	}
	cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
	uint8_t* keys_dev = scratch;
	uint8_t* heap_dev = keys_dev + keys_cap;
	uint8_t* flags_dev = heap_dev + heap_cap;
	uint8_t* const_dev = flags_dev + flags_cap;
	uint32_t* counts_dev = reinterpret_cast<uint32_t*>(const_dev + const_cap);
	uint32_t counts[kStaticMaxRowgroups];
	for (uint32_t r = 0; r < n_rg; ++r) {
		counts[r] = t->rowgroups[rg_begin + r].n_tuples;
	}
	FLS_CUDA(cudaMemsetAsync(mask_dev, 0, need, s)); // vectors nobody writes select nothing (as flsgpu_filter_mask)
	fls_static_scratch_clear_kernel<<<(uint32_t)((flags_cap / 16u + 255u) / 256u), 256, 0, s>>>(keys_cap + heap_cap, flags_cap);
	FLS_CUDA(cudaGetLastError());
	FLS_CUDA(cudaMemcpyAsync(counts_dev, counts, (size_t)n_rg * 4u, cudaMemcpyHostToDevice, s)); // pageable sources: staged before
	if (filter->str_len) {                                                                         // the call returns
		FLS_CUDA(cudaMemcpyAsync(const_dev, filter->str_data, filter->str_len, cudaMemcpyHostToDevice, s));
	}
	rc = flsgpu_decode_string_dict(t, filter->col, rg_begin, n_rg, keys_dev, keys_cap, heap_dev, heap_cap, reinterpret_cast<uint64_t>(heap_dev), s);
	if (rc) {
		return rc;
	}
	FlsStrFilterLaunch K {};
	K.rows = keys_dev;
	K.const_data = const_dev;
	K.const_len = (uint32_t)filter->str_len;
	if (filter->str_len) {
		std::memcpy(&K.const_prefix, filter->str_data, (size_t)std::min<uint64_t>(4, filter->str_len));
	}
	K.op = (uint32_t)filter->op;
	K.key_flags_out = flags_dev;
	K.n_keys = (uint32_t)n_keys;
	FLS_CUDA(fls_str_launch_key_flags(K, s));
	FlsStrLaunch L {};
	rc = string_launch_setup(t, filter->col, rg_begin, n_rg, who, &d, &L, s);
	if (rc) {
		return rc;
	}
	L.key_flags = flags_dev;
	L.n_tuples = counts_dev;
	L.rg_keep = rg_keep_dev;
	L.filter_mask = mask_dev;
	L.filter_and = 0u;
	FLS_CUDA(fls_str_launch_rows(L, s));
	g_launches.fetch_add(3, std::memory_order_relaxed);
	return FLSGPU_OK;
}

} // extern "C"
