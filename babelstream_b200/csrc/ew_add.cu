// ew_add.cu -- instantiates every compiled variant of the add kernel (see launch_ew.cuh).
#include "launch_ew.cuh"
namespace b200 { int launch_ew_add(const EwCall& call) { return launch_ew_op<OP_ADD>(call); } }
