// The 2D index / size types (GlobalElementIndex, LocalTileIndex, TileElementSize, ...) live in dlaf/common/index2d.h in
// this build; this header keeps the reference's include path (include/dlaf/matrix/index.h) working.
#pragma once
#include "dlaf/common/index2d.h"
