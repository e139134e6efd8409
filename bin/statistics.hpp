#pragma once
/* kept for source compatibility with programs that include "statistics.hpp" from bin/ (reference bin/statistics.hpp) */
#include "stencil/statistics.hpp"
