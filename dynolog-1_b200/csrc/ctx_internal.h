// What the other translation units of libhbt_b200.so may use of an hbt_ctx (defined in
// hbt_b200.cu): its device, its stream, its error string and its launch counter.
#pragma once
#include <cuda_runtime.h>

#include "../../include/hbt_b200.h"

#define HBT_INTERNAL __attribute__((visibility("hidden")))

HBT_INTERNAL int hbt_ctx_device(const hbt_ctx* ctx);
HBT_INTERNAL cudaStream_t hbt_ctx_stream(const hbt_ctx* ctx);
HBT_INTERNAL int hbt_ctx_sm_count(const hbt_ctx* ctx);
HBT_INTERNAL void hbt_ctx_count_launches(hbt_ctx* ctx, unsigned n);
// stores a printf-style message as the ctx's last error and returns `code`
HBT_INTERNAL int hbt_ctx_fail(hbt_ctx* ctx, int code, const char* fmt, ...);
