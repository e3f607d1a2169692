// Tensor-map (TMA descriptor) helpers shared by the kernels that stage bf16 matrices with cp.async.bulk.tensor.
#pragma once

#include <cuda.h>

#include "bfk_common.cuh"

namespace bfk {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled resolved through the runtime (libbfk does not link libcuda); nullptr without a driver
EncodeTiledFn get_encode_fn();
// bf16 matrix [rows, ld] row-major; box = 64 columns x box_rows, 128-byte swizzle, OOB rows read as 0
int make_tmap(CUtensorMap* m, const void* base, int64_t rows, int ld, int box_rows);
// the same with a 32-column box (64-byte rows) and the 64-byte swizzle: one attention head of a token block
int make_tmap_head(CUtensorMap* m, const void* base, int64_t rows, int ld, int box_rows);

}  // namespace bfk
