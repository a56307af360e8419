# Exact check (rational arithmetic) of the reciprocal-FMA division of the quantise kernel: q1 == RN(a / s) for fp32 a, 11-bit s.
import random, struct
from fractions import Fraction

def rn32(x: Fraction) -> Fraction:
    """round-to-nearest-even of a rational to the fp32 grid (normal range assumed)"""
    if x == 0: return Fraction(0)
    sign = -1 if x < 0 else 1
    x = abs(x)
    e = 0
    # find e with 2^23 <= x / 2^e < 2^24
    num, den = x.numerator, x.denominator
    e = num.bit_length() - den.bit_length() - 24
    while x / Fraction(2) ** e >= 2 ** 24: e += 1
    while x / Fraction(2) ** e < 2 ** 23: e -= 1
    m = x / Fraction(2) ** e
    fl = m.numerator // m.denominator
    rem = m - fl
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and (fl & 1)): fl += 1
    return sign * fl * Fraction(2) ** e

def f32_from_bits(b): return Fraction(struct.unpack('<f', struct.pack('<I', b))[0])

random.seed(1)
bad = 0
N = int(__import__("os").environ.get("N", "60000"))
for it in range(N):
    # s: bf16 or fp16 value (<= 11 significant bits), moderate exponent
    if it & 1:
        sb = (random.randrange(100, 150) << 7 | random.randrange(128)) << 16          # bf16 bits -> fp32
    else:
        h = random.randrange(1 << 10) | 1 << 10                                        # 11-bit significand
        sb = None
        s = Fraction(h) * Fraction(2) ** random.randrange(-30, 10)
    s = f32_from_bits(sb) if sb is not None else s
    if random.random() < 0.5: s = -s
    # a: arbitrary fp32 with exponent in the guarded range, some adversarial mantissas
    man = random.choice([random.randrange(1 << 23), (1 << 23) - 1, 0, 1, (1 << 22), random.randrange(1 << 23) | 1])
    ab = (random.randrange(80, 170) << 23) | man
    a = f32_from_bits(ab) * random.choice([1, -1])
    r = rn32(1 / s)
    q0 = rn32(a * r)
    e = a - s * q0
    assert rn32(e) == e, ("residual not representable", a, s)
    q1 = rn32(q0 + e * r)
    if q1 != rn32(a / s):
        bad += 1
        print("MISMATCH", float(a), float(s))
print("checked", N, "mismatches", bad)
