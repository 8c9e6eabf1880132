// Helpers shared by the C-ABI translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vda {
int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int num_sms();
bool pdl_enabled();
int encode_bf16_map(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box);
}  // namespace vda
