// clt_spec.cu -- straight-line kernels specialised to the shipped recurrences.
#include "bbb_internal.h"
#include "clt_common.cuh"

int bbb_clt_spec_threads_per_sm(const bbb_ctx *, bool, uint32_t *) { return 0; }
int bbb_clt_spec_level(bbb_ctx *, const SpecLevel &) { return 0; }
