// tcgen05 path of the GRU layer (gru_umma.cu); gru.cu dispatches here when the engine is enabled and H % 8 == 0.
#pragma once
#include "common.cuh"
namespace tt {
uint64_t gru_fwd_workspace_umma(const tt_gru_layer* a);
int gru_layer_fwd_umma(const tt_gru_layer* a, cudaStream_t s);
uint64_t gru_bwd_workspace_umma(const tt_gru_layer_grad* a);
int gru_layer_bwd_umma(const tt_gru_layer_grad* a, cudaStream_t s);
int engine_mode();            // 0 = simt-fp32, 1 = tcgen05-bf16x3
}  // namespace tt
