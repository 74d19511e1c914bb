// Instantiates the device pipelines for bls12_377_g2 (one translation unit per group keeps the
// fully unrolled limb code of each field compiling in parallel).
#include "engine_impl.cuh"

namespace b200msm {
const GroupOps *ops_bls12_377_g2() { return make_ops<Bls12377G2>(); }
}  // namespace b200msm
