// Instantiates the device pipelines for bls12_381_g1 (one translation unit per group keeps the
// fully unrolled limb code of each field compiling in parallel).
#include "engine_impl.cuh"

namespace b200msm {
const GroupOps *ops_bls12_381_g1() { return make_ops<Bls12381G1>(); }
}  // namespace b200msm
