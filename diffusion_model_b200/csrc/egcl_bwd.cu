// placeholder until the backward kernels land
#include "common.cuh"
extern "C" int64_t egnn_backward_workspace_floats(const egnn_net_t*, int32_t, int32_t, int32_t) { return 4; }
extern "C" int egnn_net_backward(const egnn_net_t*, const egnn_graph_t*, const float*, const float*, const float*,
                                 const float*, const float*, const float*, const float*, float*, float*, float*, float*,
                                 int64_t, int32_t, int32_t, void*) {
    egnn_set_error("egnn_net_backward: not built yet");
    return EGNN_ERR_UNSUPPORTED;
}
