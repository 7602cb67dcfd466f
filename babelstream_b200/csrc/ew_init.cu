// ew_init.cu -- instantiates every compiled variant of the init kernel (see launch_ew.cuh).
#include "launch_ew.cuh"
namespace b200 { int launch_ew_init(const EwCall& call) { return launch_ew_op<OP_INIT>(call); } }
