// rt_arena_writer.hpp - serialiser for the arena of rt_arena.h (format utility only: no scene
// logic).  Sections are appended as typed vectors and laid out 16-byte aligned behind the header.
// Used by the product's host loader (csrc/host/scene_flatten.cpp) and by the oracle harness that
// walks the real reference Scene (oracle/ref_harness.cpp), so both emit byte-comparable arenas.
#pragma once
#include <cstdlib>
#include <cstring>
#include <vector>
#include "rt_arena.h"

namespace rt {

class ArenaWriter {
public:
	ArenaWriter() { std::memset(sec_, 0, sizeof(sec_)); }

	template <typename T> void put(uint32_t id, const std::vector<T>& v) {
		put_raw(id, v.empty() ? nullptr : v.data(), sizeof(T), v.size());
	}
	void put_raw(uint32_t id, const void* data, uint32_t elem_size, uint64_t count) {
		Pending p;
		p.id = id;
		p.elem_size = elem_size;
		p.count = count;
		p.bytes.assign((const unsigned char*)data, (const unsigned char*)data + (size_t)elem_size * count);
		pending_.push_back(std::move(p));
	}

	// Returns a malloc'ed arena (caller frees with free()).
	void* finish(size_t* bytes_out) {
		size_t off = align_up(sizeof(rt_arena_header));
		for (auto& p : pending_) {
			rt_section& s = sec_[p.id];
			s.id = p.id;
			s.elem_size = p.elem_size;
			s.offset = off;
			s.count = p.count;
			off = align_up(off + p.bytes.size());
		}
		unsigned char* buf = (unsigned char*)std::calloc(1, off ? off : 1);
		if (!buf) return nullptr;
		rt_arena_header h;
		std::memset(&h, 0, sizeof(h));
		h.magic = RT_ARENA_MAGIC;
		h.version = RT_ARENA_VERSION;
		h.n_sections = RT_SEC_MAX;
		h.total_bytes = off;
		std::memcpy(h.sections, sec_, sizeof(sec_));
		std::memcpy(buf, &h, sizeof(h));
		for (auto& p : pending_)
			if (!p.bytes.empty()) std::memcpy(buf + sec_[p.id].offset, p.bytes.data(), p.bytes.size());
		*bytes_out = off;
		return buf;
	}

private:
	struct Pending {
		uint32_t id, elem_size;
		uint64_t count;
		std::vector<unsigned char> bytes;
	};
	static size_t align_up(size_t x) { return (x + RT_ARENA_ALIGN - 1) / RT_ARENA_ALIGN * RT_ARENA_ALIGN; }
	rt_section sec_[RT_SEC_MAX];
	std::vector<Pending> pending_;
};

// Read-side view: typed pointers into an arena.
struct ArenaView {
	const unsigned char* base = nullptr;
	const rt_arena_header* hdr = nullptr;
	explicit ArenaView(const void* arena) : base((const unsigned char*)arena), hdr((const rt_arena_header*)arena) {}
	template <typename T> const T* ptr(uint32_t id) const { return (const T*)(base + hdr->sections[id].offset); }
	uint64_t count(uint32_t id) const { return hdr->sections[id].count; }
};

} // namespace rt
