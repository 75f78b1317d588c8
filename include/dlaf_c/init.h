/* Library start-up / shut-down of the C API (reference: include/dlaf_c/init.h:27-35, src/c_api/init.cpp:21-64).
 * The pika arguments are accepted and ignored (there is no task runtime underneath: ordering is done by CUDA
 * streams and events); the dlaf arguments accept the reference's --dlaf:print-config. */
#pragma once
#include <dlaf_c/utils.h>

DLAF_EXTERN_C void dlaf_initialize(int argc_pika, const char** argv_pika, int argc_dlaf,
                                   const char** argv_dlaf) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_finalize(void) DLAF_NOEXCEPT;
