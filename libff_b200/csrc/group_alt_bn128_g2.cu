// Instantiates the device pipelines for alt_bn128_g2 (one translation unit per group keeps the
// fully unrolled limb code of each field compiling in parallel).
#include "engine_impl.cuh"

namespace b200msm {
const GroupOps *ops_alt_bn128_g2() { return make_ops<AltBn128G2>(); }
}  // namespace b200msm
