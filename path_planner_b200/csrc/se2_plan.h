// Internal: interface between the volume driver (se2_volume.cu) and the tiled fast path
// (se2_fast.cu). Not part of the ABI.
#pragma once
#include "pp_internal.h"

// returned by pp_se2_launch_fast when it leaves the whole region to the generic kernel
constexpr pp_status PP_INTERNAL_DECLINED = 1002;

// builds (or re-uses) the plan of the handle's configuration and refuses configurations whose sample offsets
// exceed what the kernels stage (PP_ERR_RANGE); call before either kernel
pp_status pp_se2_prepare_plan(pp_handle h);
pp_status pp_se2_launch_fast(pp_handle h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1);
// runs on h->aux, ordered after everything already enqueued on h->stream; the caller joins with pp_se2_join_aux
pp_status pp_se2_generic_worklist(pp_handle h, const uint32_t *d_worklist, uint32_t n_work);
pp_status pp_se2_join_aux(pp_handle h);
