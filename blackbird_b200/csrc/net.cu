// placeholder until the network kernel lands (next commit): keeps the library linkable.
#include "common.cuh"
namespace bb {
struct NetDev { int game; };
int net_create(int, const bb_net_desc*, const float*, size_t, NetDev**) { set_error("network kernel not built yet"); return BB_ESTATE; }
void net_destroy(NetDev* n) { delete n; }
int net_forward(NetDev*, const void*, int, const int*, float*, float*, cudaStream_t) { set_error("network kernel not built yet"); return BB_ESTATE; }
}  // namespace bb
