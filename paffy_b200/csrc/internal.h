// internal.h - glue between pafgpu.cu (owns pafgpu_ctx) and paf_host.cpp (PAF text layer above the C ABI). Not part of the ABI.
#pragma once
#include "../../include/paffy_b200.h"
// store a message for pafgpu_last_error; returns code
int pafgpu_internal_fail(pafgpu_ctx *ctx, int code, const char *msg);
