// Entry points whose kernels are still being written; they fail loudly (no CPU fallback).
#include "../../include/v2m_cuda.h"
#include "common.cuh"

extern "C" {

int v2m_cuda_plan_row_lengths(v2m_plan_t, int, uint64_t *) { return v2m::fail(V2M_ERR_STATE, "plan_row_lengths: not implemented yet"); }
int v2m_cuda_emit_rows_unaligned(v2m_plan_t, uint32_t, uint32_t, const uint64_t *, uint8_t *) { return v2m::fail(V2M_ERR_STATE, "emit_rows_unaligned: not implemented yet"); }
int v2m_cuda_segment_classes(v2m_graph_t, const uint64_t *, uint64_t, uint32_t *, uint32_t *, uint32_t *, uint8_t *, uint64_t *, uint32_t *, uint64_t)
{ return v2m::fail(V2M_ERR_STATE, "segment_classes: not implemented yet"); }

}
