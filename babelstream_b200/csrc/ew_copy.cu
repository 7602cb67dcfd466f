// ew_copy.cu -- instantiates every compiled variant of the copy kernel (see launch_ew.cuh).
#include "launch_ew.cuh"
namespace b200 { int launch_ew_copy(const EwCall& call) { return launch_ew_op<OP_COPY>(call); } }
