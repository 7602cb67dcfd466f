// ew_mul.cu -- instantiates every compiled variant of the mul kernel (see launch_ew.cuh).
#include "launch_ew.cuh"
namespace b200 { int launch_ew_mul(const EwCall& call) { return launch_ew_op<OP_MUL>(call); } }
