"""Synthetic MSM workloads generated with the engine itself (bench.py, examples).

Random bases are P_i = k_i * G for uniformly random k_i: a small fixed-base window table of the
group generator is built once with Python integers (affine chord-and-tangent, a few thousand
additions), then `batch_exp` + `batch_to_special` run on the GPU through the C ABI -- the same
recipe a libff user follows (get_window_table + batch_exp, multiexp.tcc:809-915), so no CPU
group arithmetic touches the n points.  Scalars are uniform in [0, r) drawn with numpy; any value
below r is a valid Montgomery representation (fp.hpp:43), so they are used as `mont_repr`
directly and their plain value is s * R^-1 mod r.

Curve constants: alt_bn128_init.cpp:43-122,287-290,358-369; bls12_377_init.cpp:58-151,301,324-331;
bw6_761_init.cpp:35-146,267-296 (all with a = 0).
"""
from __future__ import annotations

import numpy as np

from . import (ALT_BN128_G1, ALT_BN128_G2, BLS12_377_G1, BLS12_377_G2, BLS12_381_G1, BLS12_381_G2,
               BW6_761_G1, BW6_761_G2, FORM_NORMAL, batch_exp, batch_to_special)

_BN_R = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001
_BN_Q = 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47
_BLS_R = 0x12AB655E9A2CA55660B44D1E5C37B00159AA76FED00000010A11800000000001
_BLS_Q = 0x1AE3A4617C510EAC63B05C06CA1493B1A22D9F300F5138F1EF3622FBA094800170B5D44300000008508C00000000001
_BLS381_R = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
_BLS381_Q = 0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB
_BW6_Q = 0x122E824FB83CE0AD187C94004FAFF3EB926186A81D14688528275EF8087BE41707BA638E584E91903CEBAFF25B423048689C8ED12F9FD9071DCD3DC73EBFF2E98A116C25667A8F8160CF8AEEAF0A437E6913E6870000082F49D00000000008B

# group id -> (r, q, limbs64 of Fr, limbs64 of Fq, extension degree, generator)
CURVES = {
    ALT_BN128_G1: (_BN_R, _BN_Q, 4, 4, 1, (1, 2)),
    ALT_BN128_G2: (_BN_R, _BN_Q, 4, 4, 2, (
        (10857046999023057135944570762232829481370756359578518086990519993285655852781,
         11559732032986387107991004021392285783925812861821192530917403151452391805634),
        (8495653923123431417604973247489272438418190587263600148770280649306958101930,
         4082367875863433681332203403145435568316851327593401208105741076214120093531))),
    BLS12_377_G1: (_BLS_R, _BLS_Q, 4, 6, 1, (
        81937999373150964239938255573465948239988671502647976594219695644855304257327692006745978603320413799295628339695,
        241266749859715473739788878240585681733927191168601896383759122102112907357779751001206799952863815012735208165030)),
    BW6_761_G1: (_BLS_Q, _BW6_Q, 6, 12, 1, (
        6238772257594679368032145693622812838779005809760824733138787810501188623461307351759238099287535516224314149266511977132140828635950940021790489507611754366317801811090811367945064510304504157188661901055903167026722666149426237,
        2101735126520897423911504562215834951148127555913367997162789335052900271653517958562461315794228241561913734371411178226936527683203879553093934185950470971848972085321797958124416462268292467002957525517188485984766314758624099)),
    # the remaining groups of the same families (generators: bls12_377_init.cpp, bw6_761_init.cpp,
    # bls12_381_init.cpp; the values are cross-checked against the compiled reference by
    # tests/test_oracle.py through oracle/pymodel.py, which holds the same constants)
    BLS12_377_G2: (_BLS_R, _BLS_Q, 4, 6, 2, (
        (111583945774695116443911226257823823434468740249883042837745151039122196680777376765707574547389190084887628324746,
         129066980656703085518157301154335215886082112524378686555873161080604845924984124025594590925548060469686767592854),
        (168863299724668977183029941347596462608978380503965103341003918678547611204475537878680436662916294540335494194722,
         233892497287475762251335351893618429603672921469864392767514552093535653615809913098097380147379993375817193725968))),
    BW6_761_G2: (_BLS_Q, _BW6_Q, 6, 12, 1, (
        6445332910596979336035888152774071626898886139774101364933948236926875073754470830732273879639675437155036544153105017729592600560631678554299562762294743927912429096636156401171909259073181112518725201388196280039960074422214428,
        562923658089539719386922163444547387757586534741080263946953401595155211934630598999300396317104182598044793758153214972605680357108252243146746187917218885078195819486220416605630144001533548163105316661692978285266378674355041)),
    BLS12_381_G1: (_BLS381_R, _BLS381_Q, 4, 6, 1, (
        3685416753713387016781088315183077757961620795782546409894578378688607592378376318836054947676345821548104185464507,
        1339506544944476473020471379941921221584933875938349620426543736416511423956333506472724655353366534992391756441569)),
    BLS12_381_G2: (_BLS381_R, _BLS381_Q, 4, 6, 2, (
        (352701069587466618187139116011060144890029952792775240219908644239793785735715026873347600343865175952761926303160,
         3059144344244213709971259814753781636986470325476647558659373206291635324768958432433509563104347017837885763365758),
        (1985150602287291935568054521177171638300868978215655730859378665066344726373823718423869104263333984641494340347905,
         927553665492332455747201965776037880757740193453592970025027978793976877002675564980949289727957565575433344219582))),
}
# u^2 = -NON_RESIDUE in the quadratic extension of the coordinate field
NON_RESIDUE = {ALT_BN128_G2: 1, BLS12_377_G2: 5, BLS12_381_G2: 1}


class _K:
    """coordinate field: Fq (deg 1, ints) or Fq2 = Fq[u]/(u^2 + nr) (deg 2, pairs)"""

    def __init__(self, q, deg, nr=1):
        self.q, self.deg, self.nr = q, deg, nr

    def lift(self, a):
        return a if self.deg == 2 else (a, 0)

    def drop(self, a):
        return a if self.deg == 2 else a[0]

    def add(self, a, b):
        return ((a[0] + b[0]) % self.q, (a[1] + b[1]) % self.q)

    def sub(self, a, b):
        return ((a[0] - b[0]) % self.q, (a[1] - b[1]) % self.q)

    def mul(self, a, b):
        q = self.q
        return ((a[0] * b[0] - self.nr * a[1] * b[1]) % q, (a[0] * b[1] + a[1] * b[0]) % q)

    def inv(self, a):
        d = pow((a[0] * a[0] + self.nr * a[1] * a[1]) % self.q, -1, self.q)
        return ((a[0] * d) % self.q, (-a[1] * d) % self.q)


def _affine_add(K, P, Q):
    if P is None:
        return Q
    if Q is None:
        return P
    (x1, y1), (x2, y2) = P, Q
    if x1 == x2:
        if K.add(y1, y2) == (0, 0):
            return None
        lam = K.mul(K.mul((3, 0), K.mul(x1, x1)), K.inv(K.add(y1, y1)))
    else:
        lam = K.mul(K.sub(y2, y1), K.inv(K.sub(x2, x1)))
    x3 = K.sub(K.sub(K.mul(lam, lam), x1), x2)
    return (x3, K.sub(K.mul(lam, K.sub(x1, x3)), y1))


def _limbs(v: int, n: int):
    return [(v >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(n)]


def _encode_point(P, q, nq, deg):
    """affine point (lifted coordinates) or None -> X | Y | Z Montgomery limbs, Z = 1"""
    R = 1 << (64 * nq)

    def coord(c):
        out = []
        for k in range(deg):
            out += _limbs((c[k] * R) % q, nq)
        return out

    if P is None:
        return coord((0, 0)) + coord((1, 0)) + coord((0, 0))
    return coord(P[0]) + coord(P[1]) + coord((1, 0))


def generator_window_table(group: int, window: int = 8) -> np.ndarray:
    """get_window_table(scalar_bits, window, G) (multiexp.tcc:809-846) for the group generator:
    array (rows, 1 << window, words), entry [r][j] = j * 2^(r*window) * G, all in special form."""
    r, q, nr, nq, deg, gen = CURVES[group]
    K = _K(q, deg, NON_RESIDUE.get(group, 1))
    g = (K.lift(gen[0]), K.lift(gen[1]))
    bits = r.bit_length()
    rows = (bits + window - 1) // window
    table = np.zeros((rows, 1 << window, 3 * deg * nq), dtype=np.uint64)
    base = g
    for row in range(rows):
        acc = None
        for j in range(1 << window):
            table[row, j] = _encode_point(acc, q, nq, deg)
            acc = _affine_add(K, acc, base)
        base = acc  # 2^window * base
    return table


def random_scalars(group: int, n: int, seed: int) -> np.ndarray:
    """n uniform values in [0, r) as 64-bit limb rows (used as Montgomery representations)."""
    r, _, nr, _, _, _ = CURVES[group]
    rng = np.random.default_rng(seed)
    top = r >> (64 * (nr - 1))
    out = rng.integers(0, 1 << 63, size=(n, nr), dtype=np.uint64) * np.uint64(2) + \
        rng.integers(0, 2, size=(n, nr), dtype=np.uint64)
    # rejection on the top limb only (strictly below the modulus's top limb): uniform on a
    # range that is a hair short of [0, r), which is all a throughput workload needs
    bits = int(top).bit_length()
    mask = np.uint64((1 << bits) - 1)
    hi = out[:, nr - 1] & mask
    bad = hi >= np.uint64(top)
    while bad.any():
        fresh = rng.integers(0, 1 << bits, size=int(bad.sum()), dtype=np.uint64)
        hi[bad] = fresh
        bad = hi >= np.uint64(top)
    out[:, nr - 1] = hi
    return out


def scalars_plain(group: int, scalars: np.ndarray) -> list:
    """plain integer values of Montgomery-form scalar rows"""
    r, _, nr, _, _, _ = CURVES[group]
    rinv = pow(1 << (64 * nr), -1, r)
    return [(sum(int(row[j]) << (64 * j) for j in range(nr)) * rinv) % r for row in scalars]


def encode_scalars(group: int, values) -> np.ndarray:
    r, _, nr, _, _, _ = CURVES[group]
    R = 1 << (64 * nr)
    return np.array([_limbs((v % r) * R % r, nr) for v in values], dtype=np.uint64).reshape(-1, nr)


def random_bases(group: int, n: int, seed: int, table: np.ndarray = None, window: int = 8,
                 batch: int = 1 << 22):
    """n random bases k_i * G in special form (device batch_exp + batch_to_special).
    Returns (bases, k) where k are the Montgomery-form scalar rows used."""
    r, _, nr, nq, deg, _ = CURVES[group]
    if table is None:
        table = generator_window_table(group, window)
    k = random_scalars(group, n, seed)
    out = np.zeros((n, 3 * deg * nq), dtype=np.uint64)
    rows = list(table)
    for lo in range(0, n, batch):
        hi = min(n, lo + batch)
        pts = batch_exp(group, r.bit_length(), window, rows, k[lo:hi])
        out[lo:hi] = batch_to_special(group, pts)
    return out, k
