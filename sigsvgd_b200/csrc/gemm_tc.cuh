// tcgen05 / TMA tensor-core path for the two N^2 contractions (declared here, gemm_tc.cu).
#pragma once
#include "common.cuh"

namespace sigsvgd {

// bytes of scratch the tensor-core path needs inside the phi workspace
size_t tc_workspace_bytes(int64_t B, int64_t N, int64_t D, int64_t C);
// true when the tcgen05 kernels handle this problem (else the SIMT fp32 kernels run)
bool tc_path_enabled(int64_t B, int64_t N, int64_t D, int64_t C, uint32_t flags);
// split V's feature block into fp16 hi/lo planes and write dist[N,N] = pairwise sq. distances
int tc_pack_and_sqdist(const float* V, int64_t ldv, int64_t P, int64_t N, int64_t D, float* dist,
                       void* tcws, cudaStream_t st);
// out[N,C] = K[N,N] . V[N,C] with both operands split into fp16 hi/lo planes
int tc_matmul_kv(const float* K, int64_t ldk, const float* V, int64_t ldv, int64_t N, int64_t C,
                 float* out, int64_t ldo, void* tcws, cudaStream_t st);

}  // namespace sigsvgd
