// ew_nstream.cu -- instantiates every compiled variant of the nstream kernel (see launch_ew.cuh).
#include "launch_ew.cuh"
namespace b200 { int launch_ew_nstream(const EwCall& call) { return launch_ew_op<OP_NSTREAM>(call); } }
