// slab_rows.h — band-of-grid-rows ownership table shared by host and device code.
#pragma once
#include <stdint.h>
namespace fluid {
constexpr int SLAB_MAX_RANKS = 16;
struct SlabRows {
    uint32_t bounds[SLAB_MAX_RANKS + 1];  // rank r owns grid rows [bounds[r], bounds[r+1])
    int world, rank;
};
}  // namespace fluid
