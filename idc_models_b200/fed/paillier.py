"""Paillier additively homomorphic encryption — the protection mechanism the reference uses through the
``phe`` package (/root/reference/secure_fed_model.py:79,109-121). ``phe`` is not
installable here, so the arithmetic is restated on Python integers:

    public_key, private_key = generate_paillier_keypair(n_length=2048)
    c = public_key.encrypt(0.25)            # EncryptedNumber
    d = (c + public_key.encrypt(1.5)) / 2   # E(a) + E(b), scalar multiply / divide — what Server.aggregate needs
    private_key.decrypt(d)                  # 0.875

It is the *parity / oracle* path (``secure_fed_model.py --protect paillier``): a 2048-bit modular
exponentiation per weight costs milliseconds on the host, which is exactly why the product path is the
pairwise-mask ring sum on the GPU (fed/secure.py, csrc/comm.cu). Floats are encoded as fixed-point integers
with ``FRAC_BITS`` fractional bits; products of encodings carry their scale along (``EncryptedNumber.scale``),
negative values live in the upper half of Z_n like phe's encoding.
"""
from __future__ import annotations

import random

FRAC_BITS = 40

_SMALL_PRIMES = [3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37, 41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89, 97]


def _is_probable_prime(n: int, rng: random.Random, rounds: int = 24) -> bool:
    if n < 2:
        return False
    for p in [2] + _SMALL_PRIMES:
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for _ in range(rounds):
        a = rng.randrange(2, n - 1)
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def _random_prime(bits: int, rng: random.Random) -> int:
    while True:
        c = rng.getrandbits(bits) | (1 << (bits - 1)) | 1
        if _is_probable_prime(c, rng):
            return c


class EncryptedNumber:
    __slots__ = ("public_key", "c", "scale")

    def __init__(self, public_key: "PaillierPublicKey", c: int, scale: int):
        self.public_key, self.c, self.scale = public_key, c, scale

    def _rescaled(self, scale: int) -> "EncryptedNumber":
        if scale == self.scale:
            return self
        if scale < self.scale:
            raise ValueError("cannot lower the scale of a ciphertext")
        return EncryptedNumber(self.public_key, pow(self.c, 1 << (scale - self.scale), self.public_key.nsquare), scale)

    def __add__(self, other):
        pk = self.public_key
        if isinstance(other, EncryptedNumber):
            if other.public_key is not pk and other.public_key.n != pk.n:
                raise ValueError("ciphertexts under different keys")
            s = max(self.scale, other.scale)
            a, b = self._rescaled(s), other._rescaled(s)
            return EncryptedNumber(pk, a.c * b.c % pk.nsquare, s)
        if other == 0:                      # lets sum() / np.mean start from 0
            return self
        return self + pk.encrypt(float(other), scale=self.scale)

    __radd__ = __add__

    def __mul__(self, scalar):
        if isinstance(scalar, EncryptedNumber):
            raise TypeError("Paillier is additively homomorphic only")
        pk = self.public_key
        m = pk.encode(float(scalar), FRAC_BITS)
        return EncryptedNumber(pk, pow(self.c, m, pk.nsquare), self.scale + FRAC_BITS)

    __rmul__ = __mul__

    def __truediv__(self, scalar):
        return self * (1.0 / float(scalar))

    def __neg__(self):
        return self * -1.0


class PaillierPublicKey:
    def __init__(self, n: int):
        self.n = n
        self.nsquare = n * n
        self.max_int = n // 3 - 1            # phe's convention: the top third flags overflow
        self._rng = random.SystemRandom()

    def encode(self, value: float, scale: int) -> int:
        m = int(round(value * (1 << scale)))
        if abs(m) > self.max_int:
            raise OverflowError("value does not fit the plaintext space")
        return m % self.n

    def raw_encrypt(self, m: int) -> int:
        n = self.n
        while True:
            r = self._rng.randrange(1, n)
            if r % n:
                break
        return (1 + n * m) % self.nsquare * pow(r, n, self.nsquare) % self.nsquare

    def encrypt(self, value: float, scale: int = FRAC_BITS) -> EncryptedNumber:
        return EncryptedNumber(self, self.raw_encrypt(self.encode(float(value), scale)), scale)


class PaillierPrivateKey:
    def __init__(self, public_key: PaillierPublicKey, p: int, q: int):
        self.public_key = public_key
        n = public_key.n
        self.lam = (p - 1) * (q - 1)
        self.mu = pow(self.lam, -1, n)
        # Chinese-remainder decryption (what phe does too): two exponentiations with half-length exponents modulo
        # p^2 and q^2 instead of one with a full-length exponent modulo n^2 — about 3-4x less work.
        # With g = n + 1:  L_p(g^(p-1) mod p^2) = (p-1) q mod p, hence h_p = ((p-1) q)^-1 mod p (likewise for q).
        self.p, self.q = p, q
        self.psquare, self.qsquare = p * p, q * q
        self.hp = pow((p - 1) * q % p, -1, p)
        self.hq = pow((q - 1) * p % q, -1, q)
        self.p_inverse = pow(p, -1, q)

    def raw_decrypt_plain(self, c: int) -> int:
        """the textbook form: L(c^lambda mod n^2) * mu mod n (kept as the cross-check of the CRT path)"""
        n, n2 = self.public_key.n, self.public_key.nsquare
        u = pow(c, self.lam, n2)
        return (u - 1) // n * self.mu % n

    def raw_decrypt(self, c: int) -> int:
        p, q = self.p, self.q
        mp = (pow(c % self.psquare, p - 1, self.psquare) - 1) // p * self.hp % p
        mq = (pow(c % self.qsquare, q - 1, self.qsquare) - 1) // q * self.hq % q
        return mp + (mq - mp) * self.p_inverse % q * p

    def decrypt(self, enc: EncryptedNumber) -> float:
        n = self.public_key.n
        m = self.raw_decrypt(enc.c)
        if m > n - self.public_key.max_int - 1:
            m -= n
        elif m > self.public_key.max_int:
            raise OverflowError("overflow detected in decrypted number")
        return m / float(1 << enc.scale) if enc.scale < 1000 else float(m >> (enc.scale - 60)) / float(1 << 60)


def generate_paillier_keypair(n_length: int = 2048, seed: int | None = None):
    """(public_key, private_key); ``seed`` makes key generation reproducible (tests), otherwise OS entropy."""
    rng = random.Random(seed) if seed is not None else random.SystemRandom()
    while True:
        p = _random_prime(n_length // 2, rng)
        q = _random_prime(n_length - n_length // 2, rng)
        if p != q and (p * q).bit_length() == n_length:
            break
    pk = PaillierPublicKey(p * q)
    return pk, PaillierPrivateKey(pk, p, q)
