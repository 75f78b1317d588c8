/* Matrix descriptor of the C API; field-for-field the reference's (include/dlaf_c/desc.h:15-25). */
#pragma once

struct DLAF_descriptor {
  int m;    /* rows of the global matrix */
  int n;    /* columns of the global matrix */
  int mb;   /* row block size */
  int nb;   /* column block size */
  int isrc; /* process row owning the first row */
  int jsrc; /* process column owning the first column */
  int i;    /* first row of the sub-matrix (0-based): a multiple of mb (the reference requires 0) */
  int j;    /* first column of the sub-matrix (0-based): a multiple of nb (the reference requires 0) */
  int ld;   /* leading dimension of the local matrix */
};
