"""csrc/decimal.hpp (the decimal -> double conversion of the device Newick parser, Eisel-Lemire) compiled
for the host, against Python's float() - which is correctly rounded like the atof of the reference's lexer
(libpll lex_utree.l / parse_utree.y) and std::from_chars of the host parser. Every converted value must be
bit-identical; what the routine declines (status 2) goes to the host parser in the product."""
import ctypes as C
import random
import struct

import numpy as np

from linearham_b200 import _abi, host


def convert(strings):
    buf = b"\0".join(s.encode() for s in strings) + b"\0"
    n = len(strings)
    out, status, used = np.zeros(n), np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int32)
    host.lib().lhh_parse_decimal_buf(buf, n, out.ctypes.data_as(_abi._f64p), status.ctypes.data_as(_abi._i32p),
                                     used.ctypes.data_as(_abi._i32p))
    return out, status, used


def bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def check(strings, must_convert=True):
    out, status, used = convert(strings)
    n_ok = 0
    for s, v, st, u in zip(strings, out, status, used):
        if st == 0:
            assert u == len(s), (s, u)
            assert bits(v) == bits(float(s)), (s, v, float(s))
            n_ok += 1
        else:
            assert not must_convert, (s, st)
    return n_ok


def test_branch_length_shapes():
    rng = random.Random(1)
    strs = []
    for _ in range(200000):
        kind = rng.random()
        if kind < 0.4:      # repr() of a double: 15-17 significant digits (the synthetic workloads)
            strs.append(repr(rng.uniform(0, 1) * 10 ** rng.randint(-8, 1)))
        elif kind < 0.7:    # RevBayes-style fixed notation
            strs.append(f"{rng.uniform(0, 2):.{rng.randint(1, 12)}f}")
        elif kind < 0.9:    # exponents
            strs.append(f"{rng.uniform(1, 10):.{rng.randint(0, 17)}e}".replace("e", rng.choice("eE")))
        else:
            strs.append(str(rng.randint(0, 10 ** rng.randint(1, 18))))
    assert check(strs) == len(strs)


def test_random_doubles_round_trip():
    """repr(x) of random bit patterns across the whole normal range converts back to x."""
    rng = np.random.default_rng(2)
    raw = rng.integers(0, 2 ** 63, size=300000, dtype=np.int64).view(np.float64)
    raw = raw[np.isfinite(raw) & (np.abs(raw) >= 2.3e-308)]
    strs = [repr(float(x)) for x in raw] + ["-" + repr(float(x)) for x in raw[:1000]]
    assert check(strs) == len(strs)


def test_halfway_and_hard_cases():
    rng = random.Random(3)
    strs = ["9007199254740993", "9007199254740992", "9007199254740991", "9007199254740995", "1e23", "8.5e22", "2.5e-1",
            "0.1", "0.3", "1.7976931348623157e308", "2.2250738585072014e-308", "4503599627370496.5", "4503599627370497.5",
            "0.500000000000000166533453693773481063544750213623046875", "1.0000000000000002", "1.00000000000000011102230246251565404",
            "0", "0.0", "-0.0", "0e5", "000.000", "1.", ".5", "-.5", "5e-1", "5E+0", "0.000001", "1e-6", "1e22", "1e-22", "123456789012345678",
            "1234567890123456789", "12345678901234567890", "0.1234567890123456789", "1e308", "1e-307"]
    # exact ties between two doubles: (2m+1) * 2^(e-1) written out in full for small exponents
    from fractions import Fraction
    for _ in range(2000):
        m = rng.randint(2 ** 52, 2 ** 53 - 1)
        fr = (2 * m + 1) * Fraction(2) ** (rng.randint(-10, 10) - 1)
        k = fr.denominator.bit_length() - 1           # the decimal expansion of a dyadic rational is finite
        digits = str(fr.numerator * 5 ** k)
        strs.append(digits if k == 0 else (digits[:-k] or "0") + "." + digits[-k:].rjust(k, "0"))
    check(strs, must_convert=False)
    out, status, used = convert(strs)
    assert (status == 0).sum() > 0.5 * len(strs)  # ties with <= 19 digits are decided here, not declined


def test_declined_spellings():
    out, status, used = convert(["inf", "nan", "+1.5", "0x1p-3", "1e", "1e+", "-", "", "abc", "1e-400", "1e400",
                                 "4.9e-324", "12345678901234567891"])
    assert status[0] == 1 and status[1] == 1 and status[2] == 1     # not a plain number: nothing consumed
    assert status[3] == 0 and used[3] == 1                           # "0" then 'x': the caller sees a non-delimiter
    assert status[4] == 2 and status[5] == 2 and status[6] == 2
    assert status[7] == 1 and status[8] == 1
    assert all(status[9:] == 2)
