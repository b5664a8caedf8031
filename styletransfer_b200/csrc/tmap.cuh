// Cached cuTensorMapEncodeTiled wrapper (defined in conv_tc.cu): bf16 tensor, dims innermost
// first, contiguous, zero out-of-bounds fill, 128-byte or 64-byte swizzle.
#pragma once
#include <cuda.h>
#include <stdint.h>
// pixel_stride > 1: traversal (element) stride of dims 1 and 2 (W and H of an NHWC tensor); the box
// dims then span box * stride elements of which every stride-th is loaded.
int st_make_tmap(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint32_t* box, int swizzle_bytes,
                 int pixel_stride = 1);
