#pragma once
#include <cuda.h>
#include <stdint.h>
int b2c_make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box);
