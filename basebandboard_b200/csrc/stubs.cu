// stubs.cu -- entry points not implemented yet (removed one by one as the kernels land).
#include "bbb_internal.h"
extern "C" {
uint32_t bbb_prbs_tap(uint32_t k)
{
    switch (k) {   // prbs.py:14
    case 7: return 6; case 9: return 5; case 11: return 9; case 15: return 14;
    case 20: return 3; case 23: return 18; case 31: return 28; default: return 0;
    }
}
int bbb_prbs_stream(bbb_ctx *c, uint32_t, uint32_t, uint64_t, uint64_t, uint64_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_prbs_stream_dev(bbb_ctx *c, uint32_t, uint32_t, uint64_t, uint64_t, uint64_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_prbs_detect(bbb_ctx *c, uint32_t, uint64_t, uint64_t, uint32_t, uint64_t, bbb_det_state *, bbb_det_counters *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_rnghunt_eval(bbb_ctx *c, uint32_t, uint64_t, const uint64_t *, uint8_t *, uint64_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_rnghunt_candidate(bbb_ctx *c, uint32_t, uint64_t, uint64_t, uint64_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
int bbb_rnghunt_sweep(bbb_ctx *c, uint32_t, uint64_t, uint64_t, uint64_t, uint64_t, uint32_t, uint64_t *, uint32_t *, uint64_t *, uint8_t *) { return c ? c->fail(BBB_EUNSUPPORTED, "not implemented") : BBB_EINVAL; }
}
