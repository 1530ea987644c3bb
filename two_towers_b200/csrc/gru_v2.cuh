// "v2" GRU stack pipeline (gru_v2.cu): time-major fp16 planes, 3-pass forward, 1-pass loss-scaled backward.
#pragma once
#include "common.cuh"
namespace tt {
bool v2_supported(const tt_gru_stack* a);
uint64_t gru_stack_saved_bytes_v2(const tt_gru_stack* a);
uint64_t gru_stack_fwd_workspace_v2(const tt_gru_stack* a);
int gru_stack_fwd_v2(const tt_gru_stack* a, cudaStream_t s);
uint64_t gru_stack_bwd_workspace_v2(const tt_gru_stack_grad* a);
int gru_stack_bwd_v2(const tt_gru_stack_grad* a, cudaStream_t s);
int engine_mode();            // 0 = simt-fp32, 1 = tcgen05-bf16x3 (per layer), 2 = v2 stack pipeline
}  // namespace tt
