// stubs.cu -- entry points not implemented yet (removed one by one as the kernels land).
#include "bbb_internal.h"
extern "C" {
int bbb_rnghunt_eval(bbb_ctx *c, uint32_t, uint64_t, const uint64_t *, uint8_t *, uint64_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_rnghunt_candidate(bbb_ctx *c, uint32_t, uint64_t, uint64_t, uint64_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_rnghunt_sweep(bbb_ctx *c, uint32_t, uint64_t, uint64_t, uint64_t, uint64_t, uint32_t, uint64_t *, uint32_t *, uint64_t *, uint8_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
}
