// ew_triad.cu -- instantiates every compiled variant of the triad kernel (see launch_ew.cuh).
#include "launch_ew.cuh"
namespace b200 { int launch_ew_triad(const EwCall& call) { return launch_ew_op<OP_TRIAD>(call); } }
