// Thread-local lending of pinned staging memory to the ABI allocation helpers (lsr_abi.cpp).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace lsr {

struct ExportArena {
    uint8_t* base = nullptr;
    size_t cap = 0, used = 0;
    bool overflow = false;  // some alloc_component call did not fit and fell back to malloc
};
// pinned buffers whose limbs free_* must not hand to free()
void register_pinned_range(const void* base, size_t bytes);
void unregister_pinned_range(const void* base);
// alloc_component calls of THIS thread between begin and end are served from `a`
void export_arena_begin(ExportArena* a);
void export_arena_end();

}  // namespace lsr
