"""TEST INFRASTRUCTURE ONLY — pure-Python big-int model of the MSM path.

This is the slow, obviously-correct layer of the oracle: affine short-Weierstrass
arithmetic over Python integers, Montgomery-limb (de)serialisation in the reference's
in-memory layout, the reference's synthetic-input generator and the closed-form expected
MSM value for arithmetic-walk bases.  Pure-Python loops, so small cases only
(the closed form is O(n) big-int adds and is fine up to 2^26 via `walk_expected_scalar`).

Not imported by the product package (libff_b200/); only tests/, __graft_entry__.smoke() and
bench.py use it, and only as a checker.

Reference facts restated here (file:line in /root/reference):
  moduli / generators   libff/algebra/curves/alt_bn128/alt_bn128_init.cpp:43-122,287-290,358-369
                        libff/algebra/curves/bls12_377/bls12_377_init.cpp:60-151,324-331
                        libff/algebra/curves/bw6_761/bw6_761_init.cpp:38-146,285-296
  limb layout           libff/algebra/fields/bigint.hpp:28-65 (little-endian mp_limb_t data[n])
  Montgomery form       libff/algebra/fields/fp.tcc:230-235 (x*R mod p, R = 2^(64 n))
  SHA512_rng            libff/common/rng.tcc:26-71
  pippenger_optimal_c   libff/algebra/scalar_multiplication/multiexp.tcc:35-40
  log2 (ceil)           libff/common/utils.cpp:32-44
  signed digits         libff/algebra/fields/field_utils.tcc:50-100,167-203
"""
from __future__ import annotations

import hashlib
import struct
from dataclasses import dataclass

import numpy as np

# ----------------------------------------------------------------------------------------------
# Fields


@dataclass(frozen=True)
class PrimeField:
    name: str
    p: int
    limbs64: int

    @property
    def bits(self) -> int:
        return self.p.bit_length()

    @property
    def R(self) -> int:
        return 1 << (64 * self.limbs64)

    @property
    def nbytes(self) -> int:
        return 8 * self.limbs64

    def to_mont(self, x: int) -> int:
        return (x * self.R) % self.p

    def from_mont(self, x: int) -> int:
        return (x * pow(self.R, -1, self.p)) % self.p

    def inv64(self) -> int:
        return (-pow(self.p, -1, 1 << 64)) % (1 << 64)


ALT_BN128_R = PrimeField(
    "alt_bn128_Fr", 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001, 4)
ALT_BN128_Q = PrimeField(
    "alt_bn128_Fq", 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47, 4)
BLS12_377_R = PrimeField(
    "bls12_377_Fr", 0x12AB655E9A2CA55660B44D1E5C37B00159AA76FED00000010A11800000000001, 4)
BLS12_377_Q = PrimeField(
    "bls12_377_Fq",
    0x1AE3A4617C510EAC63B05C06CA1493B1A22D9F300F5138F1EF3622FBA094800170B5D44300000008508C00000000001,
    6)
BW6_761_R = PrimeField("bw6_761_Fr", BLS12_377_Q.p, 6)
BW6_761_Q = PrimeField(
    "bw6_761_Fq",
    0x122E824FB83CE0AD187C94004FAFF3EB926186A81D14688528275EF8087BE41707BA638E584E91903CEBAFF25B423048689C8ED12F9FD9071DCD3DC73EBFF2E98A116C25667A8F8160CF8AEEAF0A437E6913E6870000082F49D00000000008B,
    12)

BLS12_381_R = PrimeField(
    "bls12_381_Fr", 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001, 4)
BLS12_381_Q = PrimeField(
    "bls12_381_Fq",
    0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB,
    6)

# ----------------------------------------------------------------------------------------------
# Element arithmetic: coordinates are ints (degree 1) or 2-tuples (Fq2 = Fq[u]/(u^2 - nr)).


class Fq1Ops:
    deg = 1

    def __init__(self, field: PrimeField):
        self.f = field
        self.p = field.p

    def add(self, a, b):
        return (a + b) % self.p

    def sub(self, a, b):
        return (a - b) % self.p

    def mul(self, a, b):
        return (a * b) % self.p

    def neg(self, a):
        return (-a) % self.p

    def inv(self, a):
        return pow(a, -1, self.p)

    def zero(self):
        return 0

    def one(self):
        return 1

    def is_zero(self, a):
        return a % self.p == 0

    def from_int(self, k: int):
        return k % self.p

    def coeffs(self, a):
        return [a]

    def from_coeffs(self, c):
        return c[0]


class Fq2Ops:
    deg = 2

    def __init__(self, field: PrimeField, non_residue: int):
        self.f = field
        self.p = field.p
        self.nr = non_residue % field.p

    def add(self, a, b):
        return ((a[0] + b[0]) % self.p, (a[1] + b[1]) % self.p)

    def sub(self, a, b):
        return ((a[0] - b[0]) % self.p, (a[1] - b[1]) % self.p)

    def mul(self, a, b):
        return ((a[0] * b[0] + self.nr * a[1] * b[1]) % self.p,
                (a[0] * b[1] + a[1] * b[0]) % self.p)

    def neg(self, a):
        return ((-a[0]) % self.p, (-a[1]) % self.p)

    def inv(self, a):
        # 1/(a0 + a1 u) = (a0 - a1 u) / (a0^2 - nr a1^2)
        d = pow((a[0] * a[0] - self.nr * a[1] * a[1]) % self.p, -1, self.p)
        return ((a[0] * d) % self.p, (-a[1] * d) % self.p)

    def zero(self):
        return (0, 0)

    def one(self):
        return (1, 0)

    def is_zero(self, a):
        return a[0] % self.p == 0 and a[1] % self.p == 0

    def from_int(self, k: int):
        return (k % self.p, 0)

    def coeffs(self, a):
        return [a[0], a[1]]

    def from_coeffs(self, c):
        return (c[0], c[1])


# ----------------------------------------------------------------------------------------------
# Groups.  Affine points are (x, y) or None for the point at infinity.


@dataclass(frozen=True)
class Group:
    gid: int
    name: str
    fr: PrimeField
    fq: PrimeField
    K: object  # coordinate-field ops
    gen: tuple
    b: object  # curve coefficient (a = 0 everywhere)
    projective: bool  # bw6_761: homogeneous projective; others Jacobian

    @property
    def coord_bytes(self) -> int:
        return self.fq.nbytes * self.K.deg

    @property
    def group_bytes(self) -> int:
        return 3 * self.coord_bytes

    @property
    def scalar_bytes(self) -> int:
        return self.fr.nbytes

    # -- affine arithmetic ------------------------------------------------------------------
    def is_on_curve(self, P) -> bool:
        if P is None:
            return True
        K = self.K
        x, y = P
        return K.sub(K.mul(y, y), K.add(K.mul(K.mul(x, x), x), self.b)) == K.zero()

    def neg(self, P):
        return None if P is None else (P[0], self.K.neg(P[1]))

    def add(self, P, Q):
        K = self.K
        if P is None:
            return Q
        if Q is None:
            return P
        x1, y1 = P
        x2, y2 = Q
        if x1 == x2:
            if K.is_zero(K.add(y1, y2)):
                return None
            three = K.from_int(3)
            lam = K.mul(K.mul(three, K.mul(x1, x1)), K.inv(K.add(y1, y1)))
        else:
            lam = K.mul(K.sub(y2, y1), K.inv(K.sub(x2, x1)))
        x3 = K.sub(K.sub(K.mul(lam, lam), x1), x2)
        y3 = K.sub(K.mul(lam, K.sub(x1, x3)), y1)
        return (x3, y3)

    def mul(self, k: int, P):
        k %= self.fr.p
        R = None
        for bit in bin(k)[2:] if k else "":
            R = self.add(R, R)
            if bit == "1":
                R = self.add(R, P)
        return R

    def msm(self, scalars, points):
        acc = None
        for s, P in zip(scalars, points):
            acc = self.add(acc, self.mul(s, P))
        return acc

    # -- (de)serialisation in the reference's in-memory layout ---------------------------------
    def _coord_to_limbs(self, c) -> list:
        out = []
        for v in self.K.coeffs(c):
            m = self.fq.to_mont(v)
            out += [(m >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(self.fq.limbs64)]
        return out

    def _coord_from_limbs(self, limbs):
        n = self.fq.limbs64
        cs = []
        for k in range(self.K.deg):
            m = sum(int(limbs[k * n + i]) << (64 * i) for i in range(n))
            cs.append(self.fq.from_mont(m))
        return self.K.from_coeffs(cs)

    def encode_affine(self, P) -> np.ndarray:
        """Affine point -> X|Y|Z limbs in the reference's "special" form (Z = 1), or the
        canonical zero (0, 1, 0) (alt_bn128_init.cpp:287-288, bw6_761_init.cpp:285-286)."""
        K = self.K
        if P is None:
            coords = [K.zero(), K.one(), K.zero()]
        else:
            coords = [P[0], P[1], K.one()]
        return np.array(sum((self._coord_to_limbs(c) for c in coords), []), dtype=np.uint64)

    def encode_projective(self, P, z) -> np.ndarray:
        """Affine point lifted to the group's native non-normalised representation with the
        given non-zero Z (Jacobian: X = x z^2, Y = y z^3; bw6_761: X = x z, Y = y z)."""
        K = self.K
        if P is None:
            return self.encode_affine(None)
        if self.projective:
            coords = [K.mul(P[0], z), K.mul(P[1], z), z]
        else:
            z2 = K.mul(z, z)
            coords = [K.mul(P[0], z2), K.mul(P[1], K.mul(z2, z)), z]
        return np.array(sum((self._coord_to_limbs(c) for c in coords), []), dtype=np.uint64)

    def decode(self, limbs):
        """X|Y|Z limbs (any representation the reference accepts) -> affine point or None."""
        K = self.K
        n = self.fq.limbs64 * K.deg
        X = self._coord_from_limbs(limbs[0:n])
        Y = self._coord_from_limbs(limbs[n:2 * n])
        Z = self._coord_from_limbs(limbs[2 * n:3 * n])
        if self.projective:
            if K.is_zero(X) and K.is_zero(Z):  # bw6_761_g1.cpp:79-82
                return None
            zi = K.inv(Z)
            return (K.mul(X, zi), K.mul(Y, zi))
        if K.is_zero(Z):  # alt_bn128_g1.cpp:91
            return None
        zi = K.inv(Z)
        zi2 = K.mul(zi, zi)
        return (K.mul(X, zi2), K.mul(Y, K.mul(zi2, zi)))

    def encode_points(self, pts) -> np.ndarray:
        return np.stack([self.encode_affine(P) for P in pts]) if len(pts) else np.zeros(
            (0, self.group_bytes // 8), dtype=np.uint64)

    def encode_scalars(self, scalars) -> np.ndarray:
        n = self.fr.limbs64
        out = np.zeros((len(scalars), n), dtype=np.uint64)
        for i, s in enumerate(scalars):
            m = self.fr.to_mont(s % self.fr.p)
            for j in range(n):
                out[i, j] = (m >> (64 * j)) & 0xFFFFFFFFFFFFFFFF
        return out

    def decode_scalars(self, arr) -> list:
        n = self.fr.limbs64
        return [self.fr.from_mont(sum(int(row[j]) << (64 * j) for j in range(n))) for row in arr]


_KQ_BN = Fq1Ops(ALT_BN128_Q)
_KQ2_BN = Fq2Ops(ALT_BN128_Q, ALT_BN128_Q.p - 1)  # alt_bn128_init.cpp:138-140
_KQ_BLS = Fq1Ops(BLS12_377_Q)
_KQ_BW6 = Fq1Ops(BW6_761_Q)

ALT_BN128_G1 = Group(0, "alt_bn128_G1", ALT_BN128_R, ALT_BN128_Q, _KQ_BN, (1, 2), 3, False)

_bn_twist_b = _KQ2_BN.mul((3, 0), _KQ2_BN.inv((9, 1)))  # alt_bn128_init.cpp:267-268
ALT_BN128_G2 = Group(
    1, "alt_bn128_G2", ALT_BN128_R, ALT_BN128_Q, _KQ2_BN,
    ((10857046999023057135944570762232829481370756359578518086990519993285655852781,
      11559732032986387107991004021392285783925812861821192530917403151452391805634),
     (8495653923123431417604973247489272438418190587263600148770280649306958101930,
      4082367875863433681332203403145435568316851327593401208105741076214120093531)),
    _bn_twist_b, False)

BLS12_377_G1 = Group(
    2, "bls12_377_G1", BLS12_377_R, BLS12_377_Q, _KQ_BLS,
    (81937999373150964239938255573465948239988671502647976594219695644855304257327692006745978603320413799295628339695,
     241266749859715473739788878240585681733927191168601896383759122102112907357779751001206799952863815012735208165030),
    1, False)

BW6_761_G1 = Group(
    3, "bw6_761_G1", BW6_761_R, BW6_761_Q, _KQ_BW6,
    (6238772257594679368032145693622812838779005809760824733138787810501188623461307351759238099287535516224314149266511977132140828635950940021790489507611754366317801811090811367945064510304504157188661901055903167026722666149426237,
     2101735126520897423911504562215834951148127555913367997162789335052900271653517958562461315794228241561913734371411178226936527683203879553093934185950470971848972085321797958124416462268292467002957525517188485984766314758624099),
    BW6_761_Q.p - 1, True)

# The remaining groups of the same families; generators as the compiled
# reference reports them (bls12_377_init.cpp:392-420, bw6_761_init.cpp:340-360,
# bls12_381_init.cpp:441-450, 517-534).
_KQ2_BLS = Fq2Ops(BLS12_377_Q, BLS12_377_Q.p - 5)  # bls12_377_init.cpp:170-176
_KQ_381 = Fq1Ops(BLS12_381_Q)
_KQ2_381 = Fq2Ops(BLS12_381_Q, BLS12_381_Q.p - 1)  # bls12_381_init.cpp:137-139

BLS12_377_G2 = Group(
    4, "bls12_377_G2", BLS12_377_R, BLS12_377_Q, _KQ2_BLS,
    ((0xb997fef930828fe1b9e6a1707b8aa508a3dbfd7fe2246499c709226a0a6fef49f85b3a375363f4f8f6ea3fbd159f8a,
      0xd6ac33b84947d9845f81a57a136bfa326e915fabc8cd6a57ff133b42d00f62e4e1af460228cd5184deae976fa62596),
     (0x118dd509b2e9a13744a507d515a595dbb7e3b63df568866473790184bdf83636c94df2b7a962cb2af4337f07cb7e622,
      0x185067c6ca76d992f064a432bd9f9be832b0cac2d824d0518f77d39e76c3e146afb825f2092218d038867d7f337a010)),
    _KQ2_BLS.inv((0, 1)), False)  # y^2 = x^3 + 1/u, bls12_377_init.cpp:302-304

BW6_761_G2 = Group(
    5, "bw6_761_G2", BW6_761_R, BW6_761_Q, _KQ_BW6,
    (0x110133241d9b816c852a82e69d660f9d61053aac5a7115f4c06201013890f6d26b41c5dab3da268734ec3f1f09feb58c5bbcae9ac70e7c7963317a300e1b6bace6948cb3cd208d700e96efbc2ad54b06410cf4fe1bf995ba830c194cd025f1c,
     0x17c3357761369f8179eb10e4b6d2dc26b7cf9acec2181c81a78e2753ffe3160a1d86c80b95a59c94c97eb733293fef64f293dbd2c712b88906c170ffa823003ea96fcd504affc758aa2d3a3c5a02a591ec0594f9eac689eb70a16728c73b61),
    4, True)  # y^2 = x^3 + 4, bw6_761_init.cpp:272-282

BLS12_381_G1 = Group(
    6, "bls12_381_G1", BLS12_381_R, BLS12_381_Q, _KQ_381,
    (0x17f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55e83ff97a1aeffb3af00adb22c6bb,
     0x8b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc744a2888ae40caa232946c5e7e1),
    4, False)

BLS12_381_G2 = Group(
    7, "bls12_381_G2", BLS12_381_R, BLS12_381_Q, _KQ2_381,
    ((0x24aa2b2f08f0a91260805272dc51051c6e47ad4fa403b02b4510b647ae3d1770bac0326a805bbefd48056c8c121bdb8,
      0x13e02b6052719f607dacd3a088274f65596bd0d09920b61ab5da61bbdc7f5049334cf11213945d57e5ac7d055d042b7e),
     (0xce5d527727d6e118cc9cdc6da2e351aadfd9baa8cbdd3a76d429a695160d12c923ac9cc3baca289e193548608b82801,
      0x606c4a02ea734cc32acd2b02bc28b99cb3e287e85a763af267492ab572e99ab3f370d275cec1da1aaa9075ff05f79be)),
    (4, 4), False)  # y^2 = x^3 + 4 (1 + u), bls12_381_init.cpp:418-420

GROUPS = [ALT_BN128_G1, ALT_BN128_G2, BLS12_377_G1, BW6_761_G1,
          BLS12_377_G2, BW6_761_G2, BLS12_381_G1, BLS12_381_G2]
GROUP_BY_NAME = {g.name: g for g in GROUPS}

# ----------------------------------------------------------------------------------------------
# Synthetic inputs


def sha512_rng(field: PrimeField, idx: int) -> int:
    """libff::SHA512_rng<FieldT>(idx), as a plain integer in [0, p) (rng.tcc:26-71)."""
    nbytes = 8 * field.limbs64
    mask = (1 << field.bits) - 1
    it = 0
    while True:
        h = hashlib.sha512(struct.pack("<QQ", idx, it)).digest()
        v = int.from_bytes(h[:nbytes], "little") & mask
        if v < field.p:
            return v
        it += 1


WALK_A_IDX = 1 << 63
WALK_D_IDX = (1 << 63) + 1


def walk_params(group: Group):
    return sha512_rng(group.fr, WALK_A_IDX), sha512_rng(group.fr, WALK_D_IDX)


def walk_point(group: Group, i: int):
    a, d = walk_params(group)
    return group.mul((a + i * d) % group.fr.p, group.gen)


def walk_expected_scalar(group: Group, scalars_plain, start: int = 0) -> int:
    """k such that sum_i s_i * P_{start+i} = k * G for arithmetic-walk bases."""
    a, d = walk_params(group)
    r = group.fr.p
    s0 = 0
    s1 = 0
    for i, s in enumerate(scalars_plain):
        s0 += s
        s1 += (start + i) * s
    return (a * s0 + d * s1) % r


def weighted_scalar_sums(group: Group, scalars_mont: np.ndarray, start: int = 0):
    """(sum_i s_i, sum_i (start + i) s_i) mod r for Montgomery-form scalar rows (numpy uint64,
    n x limbs), vectorised so that 2^26 scalars take seconds: the Montgomery factor R comes out
    of the sums, the 64-bit limbs are split into 16-bit pieces and the index-weighted sums are
    taken over runs of 2^16 rows, which keeps every partial sum below 2^64."""
    r = group.fr.p
    a = np.ascontiguousarray(scalars_mont, dtype=np.uint64)
    n, limbs = a.shape
    s0 = 0
    s1 = 0
    RUN = 1 << 16
    for lo in range(0, n, RUN):
        blk = a[lo:lo + RUN]
        idx = np.arange(blk.shape[0], dtype=np.uint64)  # < 2^16
        b0 = 0
        b1 = 0
        for j in range(limbs):
            col = blk[:, j]
            for q in range(4):
                piece = (col >> np.uint64(16 * q)) & np.uint64(0xFFFF)   # < 2^16
                sh = 64 * j + 16 * q
                b0 += int(piece.sum(dtype=np.uint64)) << sh              # < 2^32
                b1 += int((piece * idx).sum(dtype=np.uint64)) << sh      # < 2^48
        s0 += b0
        s1 += b1 + (start + lo) * b0
    rinv = pow(1 << (64 * limbs), -1, r)
    return (s0 * rinv) % r, (s1 * rinv) % r


def walk_expected_scalar_fast(group: Group, scalars_mont: np.ndarray, start: int = 0) -> int:
    """walk_expected_scalar on the Montgomery-form limb rows directly (any n up to 2^32)."""
    a, d = walk_params(group)
    s0, s1 = weighted_scalar_sums(group, scalars_mont, start)
    return (a * s0 + d * s1) % group.fr.p


# ----------------------------------------------------------------------------------------------
# Window-size and digit helpers (restating the reference's integer arithmetic)


def ceil_log2(n: int) -> int:
    """libff::log2 (common/utils.cpp:32-44): ceil(log2 n), 0 for n in {0, 1}."""
    r = 1 if (n & (n - 1)) else 0
    while n > 1:
        n >>= 1
        r += 1
    return r


def pippenger_optimal_c(n: int) -> int:
    """multiexp.tcc:35-40, with size_t wrap-around for tiny n (n = 0, 1 -> c = 2)."""
    l2 = ceil_log2(n)
    return (l2 - ((l2 // 3 - 2) % (1 << 64))) % (1 << 64)


def get_digit(v: int, c: int, idx: int) -> int:
    return (v >> (c * idx)) & ((1 << c) - 1)


def signed_digits(v: int, c: int, num_digits: int) -> list:
    """field_get_signed_digits (field_utils.tcc:205-239)."""
    out = []
    carry = 0
    overflow = 0
    for i in range(num_digits):
        carry = overflow | carry
        digit = get_digit(v, c, i) + carry
        overflow = (digit >> c) & 1
        carry = (digit >> (c - 1)) & 1
        out.append((1 - overflow) * (digit - carry * (1 << c)))
    return out


# ----------------------------------------------------------------------------------------------
# Wire format of the reference's ffi layer (ffi/ffi_serialization.hpp:12-16, .tcc:98-187):
# big-endian, left-padded to the in-memory width, plain values; Fq2 as c1 | c0; points affine
# X | Y with (0, 1) for infinity.

def wire_coord(group: Group, c) -> bytes:
    return b"".join(int(v).to_bytes(group.fq.nbytes, "big") for v in group.K.coeffs(c)[::-1])


def wire_point(group: Group, P) -> bytes:
    K = group.K
    if P is None:
        return wire_coord(group, K.zero()) + wire_coord(group, K.one())
    return wire_coord(group, P[0]) + wire_coord(group, P[1])


def stream_point(group: Group, P, montgomery: bool) -> bytes:
    """libff's binary stream encoding of one element, uncompressed (curve_serialization.tcc:79-101,
    field_serialization.tcc:127-147, 197-222): X then Y, extension coefficients lowest first, every
    base-field value big-endian, either plain or its Montgomery representation; zero = (0, 1)."""
    K, f = group.K, group.fq
    x, y = (K.zero(), K.one()) if P is None else P
    conv = f.to_mont if montgomery else (lambda v: v)
    return b"".join(conv(c).to_bytes(f.nbytes, "big") for c in K.coeffs(x) + K.coeffs(y))


def wire_scalar(group: Group, s: int) -> bytes:
    return int(s).to_bytes(group.fr.nbytes, "big")


def sqrt_mod(a: int, p: int):
    """Tonelli-Shanks; None if a is not a square."""
    a %= p
    if a == 0:
        return 0
    if pow(a, (p - 1) // 2, p) != 1:
        return None
    if p % 4 == 3:
        return pow(a, (p + 1) // 4, p)
    q, s = p - 1, 0
    while q % 2 == 0:
        q //= 2
        s += 1
    z = 2
    while pow(z, (p - 1) // 2, p) != p - 1:
        z += 1
    m, c, t, r = s, pow(z, q, p), pow(a, q, p), pow(a, (q + 1) // 2, p)
    while t != 1:
        i, t2 = 0, t
        while t2 != 1:
            t2 = t2 * t2 % p
            i += 1
        b = pow(c, 1 << (m - i - 1), p)
        m, c = i, b * b % p
        t, r = t * c % p, r * b % p
    return r
