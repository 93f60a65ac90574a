// w4a16_prefill.cu -- placeholder until the tcgen05 kernel lands (next commit).
#include "w4a16_internal.cuh"
namespace w4 {
bool prefill_available() { return false; }
int launch_prefill(const Linear &, const __nv_bfloat16 *, __nv_bfloat16 *, int64_t, bool, cudaStream_t) {
    return fail("prefill kernel not built");
}
}  // namespace w4
