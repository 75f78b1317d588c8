// matrix::MatrixMirror lives in dlaf/matrix/matrix.h in this build; this header keeps the reference's include path
// (include/dlaf/matrix/matrix_mirror.h) working.
#pragma once
#include "dlaf/matrix/matrix.h"
