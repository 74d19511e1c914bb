// One Montgomery product spread over T lanes of a warp (T = 8 lanes of 3 limbs for the 24-limb
// bw6_761 field, 4 lanes of 3 limbs for the 12-limb bls12 fields); a warp runs 32 / T products
// at once.
//
// Why: the Horner over windows (k_final) is ONE chain of dependent doublings -- 378 of them for
// bw6_761 -- and a single lane walks a 24-limb product in about 6 microseconds (1152 dependent
// multiply-adds).  Splitting the N x N limb products row-wise over T lanes shortens the chain of
// one product to N iterations of 2 N / T multiply-adds plus two shuffles.
//
// Algorithm (operand scanning, same arithmetic as Fp::mul, fp.tcc:50-228): lane j of a group owns
// limbs [j S, (j + 1) S) of b and of the modulus and a window r[0 .. S+1] of the accumulator whose
// two top words overlap the two lowest of lane j + 1.  Per limb a_i:  r += a_i b_j;  lane 0
// derives q = r[0] * INV and broadcasts it;  r += q p_j;  the accumulator moves down one limb --
// every lane hands its r[0] to the lane below.  Carries between lanes stay in the overlap words
// and are resolved once, after the last row, through shared memory.
#pragma once
#include "fp.cuh"

namespace b200msm {

template <class F>
struct CoopField {
    static constexpr bool OK = false;
    static constexpr int WIN_WORDS = 1;
};

template <class P, int CM>
struct CoopField<Fp<P, CM>> {
    using F = Fp<P, CM>;
    static constexpr int N = P::N;
    static constexpr bool OK = N >= 12 && N % 4 == 0 && F::SPARE2;
    static constexpr int T = (N % 8 == 0 && N >= 16) ? 8 : 4;  // lanes per product
    static constexpr int S = N / T;                            // limbs per lane
    static constexpr int WIN_WORDS = 32 * (S + 1);             // shared scratch, 32-bit words

    // All 32 lanes call this; the lanes of one group pass identical operands and all receive
    // the group's product a b R^-1 mod p (canonical).
    B200_DEV static F mul(const F &a, const F &b, uint32_t *win, uint32_t lane) {
        const uint32_t j = lane % T, gbase = lane - j;
        uint32_t bs[S], ps[S];
#pragma unroll
        for (int k = 0; k < S; k++) {
            bs[k] = 0, ps[k] = 0;
#pragma unroll
            for (int jj = 0; jj < T; jj++)
                if ((uint32_t)jj == j) bs[k] = b.v[jj * S + k], ps[k] = P::mod(jj * S + k);
        }
        uint32_t r[S + 2];
#pragma unroll
        for (int k = 0; k < S + 2; k++) r[k] = 0;
#pragma unroll
        for (int i = 0; i < N; i++) {
            const uint32_t ai = a.v[i];
            uint64_t cy = 0, t;
#pragma unroll
            for (int k = 0; k < S; k++) {
                t = (uint64_t)ai * bs[k] + r[k] + cy;
                r[k] = (uint32_t)t, cy = t >> 32;
            }
            t = (uint64_t)r[S] + cy;
            r[S] = (uint32_t)t, r[S + 1] += (uint32_t)(t >> 32);
            const uint32_t q = __shfl_sync(0xffffffffu, r[0] * P::INV, gbase);
            cy = 0;
#pragma unroll
            for (int k = 0; k < S; k++) {
                t = (uint64_t)q * ps[k] + r[k] + cy;
                r[k] = (uint32_t)t, cy = t >> 32;
            }
            t = (uint64_t)r[S] + cy;
            r[S] = (uint32_t)t, r[S + 1] += (uint32_t)(t >> 32);
            // divide by 2^32: lane 0's r[0] is zero now, every other r[0] moves to the lane below
            uint32_t up = __shfl_down_sync(0xffffffffu, r[0], 1);
            if (j == T - 1) up = 0;
#pragma unroll
            for (int k = 0; k + 1 < S; k++) r[k] = r[k + 1];
            t = (uint64_t)r[S] + up;
            r[S - 1] = (uint32_t)t;
            r[S] = r[S + 1] + (uint32_t)(t >> 32);
            r[S + 1] = 0;
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k <= S; k++) win[lane * (S + 1) + k] = r[k];
        __syncwarp();
        // resolve the overlaps: limb j S + k = window j word k (+ window j - 1 word S when k = 0)
        F out;
        uint64_t carry = 0;
#pragma unroll
        for (int L = 0; L < N; L++) {
            const int jj = L / S, k = L % S;
            uint64_t sum = carry + win[(gbase + jj) * (S + 1) + k];
            if (k == 0 && jj > 0) sum += win[(gbase + jj - 1) * (S + 1) + S];
            out.v[L] = (uint32_t)sum;
            carry = sum >> 32;
        }
        __syncwarp();
        return F::reduce_once(out);  // a b + q p < 2 p R: nothing above limb N - 1
    }
};

}  // namespace b200msm
