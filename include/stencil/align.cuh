#pragma once
/* next_align_of(x, a): round x up to a multiple of the power-of-two a (host + device).
   Parity: reference include/stencil/align.cuh:7-9. */
#include <cstddef>
#include <cstdint>

#include "stencil/dim3.hpp"

STENCIL_HD constexpr inline size_t next_align_of(size_t x, size_t a) { return (x + a - 1) & ~(a - 1); }
/* round up to a multiple of any positive m (not necessarily a power of two) */
STENCIL_HD constexpr inline size_t round_up_multiple(size_t x, size_t m) { return (x + m - 1) / m * m; }
