// Internal (not C ABI) batch launchers of nconv_tc.cu used by the DNET pipeline: all N8 ring fills and all weight packs
// of one forward as one launch each.
#pragma once
#include <cuda_runtime.h>

namespace nconv {

struct N8RingJobs { void* buf[12]; int H[12], W[12]; int B, n; };
struct NtPackJobs { const float* w[8]; void* wpack[8]; int K[8], nsrc[8], a_first[8]; int n; };

int n8_zero_rings_multi(const N8RingJobs& jobs, cudaStream_t st);
int nconv_tc_pack_multi(const NtPackJobs& jobs, cudaStream_t st);

}  // namespace nconv
