// placeholder until nn_tc.cu lands
#include "nn.cuh"
namespace fvad {
std::unique_ptr<NnBackend> make_nn_tc(const HostWeights &, bool) { throw Error(FVAD_ERR_INVALID_ARGUMENT, "tensor-core NN backend not built"); }
}
