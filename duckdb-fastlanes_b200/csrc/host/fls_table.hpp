// fls_table.hpp — host view of an opened container (replaces TableMetadata + fastlanes::TableReader,
// reference src/reader/table_metadata.cpp:13-35).
#pragma once
#include <cstdint>
#include <memory>
#include <utility>
#include <vector>
#include "fls_format.h"

struct FlsDeviceState; // defined by the CUDA side (flsgpu_device.cu)

// File images are read into memory that is NOT zeroed first: value-initialising a 2 GB vector costs as much as reading it.
template <class T>
struct FlsNoInitAllocator : std::allocator<T> {
	template <class U>
	struct rebind {
		using other = FlsNoInitAllocator<U>;
	};
	template <class U>
	void construct(U* p) noexcept {
		::new (static_cast<void*>(p)) U;
	}
	template <class U, class... Args>
	void construct(U* p, Args&&... args) {
		::new (static_cast<void*>(p)) U(std::forward<Args>(args)...);
	}
};
using FlsImage = std::vector<uint8_t, FlsNoInitAllocator<uint8_t>>;

struct flsgpu_table {
	const uint8_t* bytes = nullptr; // whole file image
	uint64_t size = 0;
	std::shared_ptr<FlsImage> owned; // backing store when the table owns a copy (shared by flsgpu_table_share handles)
	int device_index = 0;           // which of flsgpu_init's devices this handle stages to and decodes on (flsgpu_table_bind_device)
	void* pinned = nullptr;         // cudaHostAlloc'ed copy used for H2D staging (lazily created)
	fls_file_header header {};
	const fls_column_desc* columns = nullptr;
	const fls_rowgroup_desc* rowgroups = nullptr;
	const fls_chunk_desc* chunks = nullptr; // [rg * n_columns + col]
	uint64_t n_vectors = 0;
	FlsDeviceState* dev = nullptr;

	const fls_chunk_desc& chunk(uint32_t rg, uint32_t col) const {
		return chunks[(uint64_t)rg * header.n_columns + col];
	}
};

// Validates header, footer bounds, every chunk/segment extent and every vector entry point.
int fls_table_parse(flsgpu_table* t);
void fls_device_release(flsgpu_table* t); // CUDA side; no-op when nothing is staged
