#pragma once
#include <cuda.h>
#include <stdint.h>
int b2c_make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box);
// rank-R fp32 tensor map, no swizzle, zero OOB fill (dims/box innermost first; strides in BYTES for dims 1..R-1)
int b2c_make_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box);
