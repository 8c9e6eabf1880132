// Helpers shared by the C-ABI translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vda {
int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int num_sms();  // SM count of the CURRENT device (vda_init(device) must have run for it)
bool pdl_enabled();

// Kernel attributes (cudaFuncAttributeMaxDynamicSharedMemorySize) are per device, so "configured" flags are
// kept per device: `static DeviceOnce once; int& c = once.slot(); if (c < need) { ...set attribute...; c = need; }`
constexpr int kMaxDevices = 64;
int current_device();
struct DeviceOnce {
  int level[kMaxDevices] = {};
  int& slot() { return level[current_device() & (kMaxDevices - 1)]; }
};
int encode_bf16_map(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box);
int encode_bf16_map_plain(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                          const uint32_t* box);
}  // namespace vda
