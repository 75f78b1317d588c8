// Umbrella header of the solvers (reference: include/dlaf/solver.h).
#pragma once
#include "dlaf/solver/triangular.h"
