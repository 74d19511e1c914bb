// Instantiates the device pipelines for bw6_761_g2 (one translation unit per group keeps the
// fully unrolled limb code of each field compiling in parallel).
#include "engine_impl.cuh"

namespace b200msm {
const GroupOps *ops_bw6_761_g2() { return make_ops<Bw6761G2>(); }
}  // namespace b200msm
