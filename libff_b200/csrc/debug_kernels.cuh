// Element-wise field and group operations exposed for parity tests of the leaves of the path
// (SURVEY.md §8a rows a7-a13): one thread per element, reference layout in and out.
#pragma once
#include "curves.cuh"

namespace b200msm {

enum DebugOp {
    DBG_FQ_ADD = 0,
    DBG_FQ_SUB = 1,
    DBG_FQ_MUL = 2,
    DBG_FQ_SQR = 3,
    DBG_FQ_NEG = 4,
    DBG_FQ_INV = 5,
    DBG_FR_FROM_MONT = 6,
    DBG_FR_MUL = 7,
    // 8 and 9 were the FP64-pipe product (measured and rejected; now tools/fp64_product/)
    DBG_G_ADD = 10,
    DBG_G_DBL = 11,
    DBG_G_MADD = 12,   // b must be special (Z = 1) or zero
    DBG_G_NEG = 13,
    DBG_FQ_MUL2 = 14,  // a b + (a + b)(a - b) through the shared-reduction product
    // the other code shapes of the same product (fp.cuh CALL_MIN policies): bit for bit equal
    DBG_FQ_MUL_ROLLED = 15,
    DBG_FQ_MUL2_ROLLED = 16,
    DBG_FQ_MUL_CALLED = 17,
    DBG_FQ_SQR_CALLED = 18,
    // plain integer product + separate Montgomery reduction (the halves of the lazy Fq2 product),
    // on the base prime field of the group
    DBG_FQ_MUL_SEPARATED = 19,
};

template <class C>
__global__ void k_debug_op(int op, const uint32_t *__restrict__ a, const uint32_t *__restrict__ b,
                           uint32_t *__restrict__ out, uint32_t n) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using Fr = typename C::Fr;
    using X = XYZZ<F>;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (op == DBG_FQ_MUL_SEPARATED) {
        using B = typename BaseField<F>::type;  // N words per element of the base prime field
        constexpr int per = F::N / B::N;        // 1, or 2 for a quadratic extension: both halves
        for (int h = 0; h < per; h++) {
            const B x = B::load(a + (size_t)i * F::N + h * B::N), y = B::load(b + (size_t)i * F::N + h * B::N);
            B::mul_separated(x, y).store(out + (size_t)i * F::N + h * B::N);
        }
    } else if (op >= DBG_FQ_MUL_ROLLED && op <= DBG_FQ_SQR_CALLED) {
        using FR = typename Rebind<F, -1>::type;  // every product rolled
        using FC = typename Rebind<F, 1>::type;   // every product a call
        if (op == DBG_FQ_MUL_ROLLED || op == DBG_FQ_MUL2_ROLLED) {
            const FR x = FR::load(a + (size_t)i * F::N), y = FR::load(b + (size_t)i * F::N);
            const FR r = op == DBG_FQ_MUL_ROLLED ? FR::mul(x, y) : FR::mul2(x, y, FR::add(x, y), FR::sub(x, y));
            r.store(out + (size_t)i * F::N);
        } else {
            const FC x = FC::load(a + (size_t)i * F::N);
            const FC r = op == DBG_FQ_MUL_CALLED ? FC::mul(x, FC::load(b + (size_t)i * F::N)) : FC::sqr(x);
            r.store(out + (size_t)i * F::N);
        }
    } else if (op <= DBG_FQ_INV || op == DBG_FQ_MUL2) {
        F x = F::load(a + (size_t)i * F::N), y = F::zero(), r;
        if (op <= DBG_FQ_MUL || op == DBG_FQ_MUL2) y = F::load(b + (size_t)i * F::N);
        switch (op) {
            case DBG_FQ_ADD: r = F::add(x, y); break;
            case DBG_FQ_SUB: r = F::sub(x, y); break;
            case DBG_FQ_MUL: r = F::mul(x, y); break;
            case DBG_FQ_MUL2: r = F::mul2(x, y, F::add(x, y), F::sub(x, y)); break;
            case DBG_FQ_SQR: r = F::sqr(x); break;
            case DBG_FQ_NEG: r = F::neg(x); break;
            default: r = F::inv(x); break;
        }
        r.store(out + (size_t)i * F::N);
    } else if (op == DBG_FR_FROM_MONT || op == DBG_FR_MUL) {
        Fr x = Fr::load(a + (size_t)i * Fr::N);
        Fr r = op == DBG_FR_MUL ? Fr::mul(x, Fr::load(b + (size_t)i * Fr::N)) : Fr::from_mont(x);
        r.store(out + (size_t)i * Fr::N);
    } else {
        F X1, Y1, Z1;
        R::load(a + (size_t)i * R::PW, X1, Y1, Z1);
        X p = R::to_xyzz(X1, Y1, Z1);
        if (op == DBG_G_ADD) {
            F X2, Y2, Z2;
            R::load(b + (size_t)i * R::PW, X2, Y2, Z2);
            X q = R::to_xyzz(X2, Y2, Z2);
            X::add(p, q);
        } else if (op == DBG_G_DBL) {
            p = X::dbl(p);
        } else if (op == DBG_G_MADD) {
            F X2, Y2, Z2;
            R::load(b + (size_t)i * R::PW, X2, Y2, Z2);
            if (!R::is_inf(X2, Z2)) X::madd(p, X2, Y2);
        } else {
            p = X::neg(p);
        }
        F Xo, Yo, Zo;
        R::from_xyzz(p, Xo, Yo, Zo);
        R::store(out + (size_t)i * R::PW, Xo, Yo, Zo);
    }
}

}  // namespace b200msm
