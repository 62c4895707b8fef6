// nnchain.cu -- placeholder until the merge loop lands
#include "rthc_internal.cuh"
int rthc_prim_exact(rthc_ctx *ctx, std::vector<int32_t> &, std::vector<int32_t> &, std::vector<double> &)
{
    return rthc_fail(ctx, RTHC_E_STATE, "Prim-exact single linkage not built yet");
}
int rthc_nn_chain(rthc_ctx *ctx, int, std::vector<int32_t> &, std::vector<int32_t> &, std::vector<double> &)
{
    return rthc_fail(ctx, RTHC_E_STATE, "NN-chain merge loop not built yet");
}
extern "C" int rthc_set_condensed_f64(rthc_ctx *ctx, const double *, int64_t)
{
    return rthc_fail(ctx, RTHC_E_STATE, "external condensed matrices not built yet");
}
