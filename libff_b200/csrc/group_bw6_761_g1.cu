// Instantiates the device pipelines for bw6_761_g1 (one translation unit per group keeps the
// fully unrolled limb code of each field compiling in parallel).
#include "engine_impl.cuh"

namespace b200msm {
const GroupOps *ops_bw6_761_g1() { return make_ops<Bw6761G1>(); }
}  // namespace b200msm
