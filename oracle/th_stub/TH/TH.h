/* Minimal stand-in for PyTorch-0.4's <TH/TH.h>, written for the oracle build only
 * (oracle/build_ref.py).  It lets the reference's libs/layers/roi_align_tf/src/crop_and_resize.c
 * compile UNMODIFIED: a dense tensor is a data pointer plus up to four sizes, and the five
 * accessors that file calls are provided.  resize4d only records the sizes: the harness
 * (oracle/ref_loader.py) always passes a buffer that is already large enough. */
#ifndef DETBOX_TH_STUB_H
#define DETBOX_TH_STUB_H
#include <stdlib.h>
#include <string.h>

typedef struct { float *data; long size[4]; int nDimension; } THFloatTensor;
typedef struct { int   *data; long size[4]; int nDimension; } THIntTensor;

static inline long   THFloatTensor_size(const THFloatTensor *t, int d) { return t->size[d]; }
static inline float *THFloatTensor_data(const THFloatTensor *t)        { return t->data; }
static inline int   *THIntTensor_data(const THIntTensor *t)            { return t->data; }
static inline void   THFloatTensor_resize4d(THFloatTensor *t, long a, long b, long c, long d)
{ t->size[0] = a; t->size[1] = b; t->size[2] = c; t->size[3] = d; t->nDimension = 4; }
static inline void   THFloatTensor_zero(THFloatTensor *t)
{ long n = 1; for (int i = 0; i < t->nDimension; ++i) n *= t->size[i];
  memset(t->data, 0, (size_t)n * sizeof(float)); }
#endif
