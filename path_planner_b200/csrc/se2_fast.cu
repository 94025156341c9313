// K1 fast path (tiled, shared-memory brick tables). Placeholder: declines everything so that the
// exact generic kernel runs; filled in once the generic path is parity-green.
#include "se2_plan.h"

struct Se2Plan
{
    int unused;
};

void pp_se2_plan_free(pp_handle h)
{
    delete h->plan;
    h->plan = nullptr;
}

pp_status pp_se2_launch_fast(pp_handle, uint32_t, uint32_t, uint32_t, uint32_t) { return PP_INTERNAL_DECLINED; }
