// tcgen05 / TMEM / TMA path - placeholder until the fused kernel lands (reports "unsupported").
#include "tc.cuh"

namespace b200km {
bool tc_supported(const float*, int, int) { return false; }
int tc_max_d() { return 0; }
int tc_max_k() { return 0; }
size_t tc_workspace_bytes(int, int) { return 0; }
int tc_pass(TcWorkspace*, const float*, int64_t, int, int, const float*, const float*, int32_t*, float*, double*,
            IterState*, bool, cudaStream_t) {
    return 1;
}
const char* tc_last_error() { return "tcgen05 path not built"; }
void tc_free(TcWorkspace*) {}
}  // namespace b200km
