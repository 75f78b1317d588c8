// Umbrella header of the factorizations on this path.
#pragma once
#include "dlaf/factorization/cholesky.h"
