// ctx_internal.h -- seam between the device context (mrisim.cu) and the multi-device group
// (group.cpp).  Not part of the C ABI.
//
// A submitted sprint has three phases.  mrisim_submit_sprint() runs them back to back for one
// context; a single-process group over several devices must interleave them, because NCCL wants the
// all-reduce calls of all its devices inside ONE ncclGroupStart/ncclGroupEnd (and defers them to the
// GroupEnd, so nothing that has to follow the all-reduce on the side stream may be enqueued before it).
#pragma once
#include "mrisim.h"

// phase 1: the sprint's kernels on `compute_stream` (nullptr: the context's own) into the next ring
// slot; the side stream is made to wait for them
int mrisim_ctx_phase_compute(mrisim_ctx* ctx, int32_t steps, void* compute_stream);
// phase 2: in-place SUM all-reduce of the slot on the side stream (no-op without a communicator)
int mrisim_ctx_phase_reduce(mrisim_ctx* ctx);
// phase 3: D2H of the slot and of the error flag on the side stream, completion event
int mrisim_ctx_phase_finish(mrisim_ctx* ctx);
