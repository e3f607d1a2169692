// placeholder until the tcgen05 path lands
#include "bfk_common.cuh"
namespace bfk {
int tensor_path_supported(const bfk_index*, int64_t, int) { return 0; }
size_t tensor_workspace_bytes(const bfk_index*, int64_t, int) { return 0; }
int search_tensor(bfk_index*, const float*, int64_t, int, float*, int64_t*, int64_t, cudaStream_t, bfk_search_stats*) {
    return fail("tensor path not built");
}
}
