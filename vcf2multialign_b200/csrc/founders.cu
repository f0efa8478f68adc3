// K3 segment_classes — placeholder until the kernels land; fails loudly (no CPU fallback).
#include "../../include/v2m_cuda.h"
#include "common.cuh"

extern "C" int v2m_cuda_segment_classes(v2m_graph_t, const uint64_t *, uint64_t, uint32_t *, uint32_t *, uint32_t *, uint8_t *, uint64_t *, uint32_t *, uint64_t)
{ return v2m::fail(V2M_ERR_STATE, "segment_classes: not implemented yet"); }
