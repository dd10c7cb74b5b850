"""The decimal -> binary64 arithmetic of csrc/strtod.cuh (the GPU's ``float(val[1])``,
gpu_util_stats.py:48,:74), compiled as plain C++ by g++ and held to Python's float() bit for bit -
including the grammar (underscores, inf/nan, blanks) and exact ties of 700-digit tokens.
CPU-only check of the arithmetic; the GPU tests run the same routine inside the parse kernels."""
import random
import struct
import subprocess
from fractions import Fraction
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope='module')
def checker(tmp_path_factory):
    exe = tmp_path_factory.mktemp('strtod') / 'strtod_check'
    subprocess.run(['g++', '-O2', '-std=c++17', '-o', str(exe), str(ROOT / 'tests/host/strtod_check.cpp')], check=True)

    def run(tokens):
        out = subprocess.run([str(exe)], input='\n'.join(tokens) + '\n', capture_output=True, text=True, check=True)
        res = []
        run.fast = []
        for line in out.stdout.splitlines():
            ok, bits, used, fbits = line.split()
            res.append((int(ok), int(bits, 16)))
            run.fast.append((int(used), int(fbits, 16)))
        return res
    return run


def expected(token):
    """float() of Python 2, the reference's interpreter: Python 3's float() minus the underscores that
    3.6 added (everything else about the grammar and the rounding is the same since 2.7)."""
    if '_' in token:
        return (0, 0)
    try:
        v = float(token)
    except ValueError:
        return (0, 0)
    bits = struct.unpack('<Q', struct.pack('<d', v))[0]
    return (1, bits)


def same(got, want):
    if got == want:
        return True
    nan = lambda b: (b & 0x7FF0000000000000) == 0x7FF0000000000000 and (b & 0xFFFFFFFFFFFFF)
    return got[0] == want[0] == 1 and nan(got[1]) and nan(want[1])     # any NaN is a NaN


def exact_decimal(fr):
    n, d = fr.numerator, fr.denominator
    s, r = str(n // d) + '.', n % d
    while r:
        r *= 10
        s += str(r // d)
        r %= d
    return s


GRAMMAR = ['0', '1', '-1', '+1.5', '1e5', '1E5', '.5', '5.', '1.e3', '.5e-3', '1_000', '1_0.0_1e1_0', 'inf', '-Inf',
           'INFINITY', 'nan', '-nan', 'NaN', ' 1.5 ', '\t2\r', '1e', 'e5', '.', '', ' ', '1__0', '_1', '1_', '1_.5', '1._5',
           '1e_5', '1e5_', '0x10', 'infinit', '1.2.3', '1e5.0', '--1', '+', '-', '1 2', '1e+', '-.', '+.e1', '00012.50',
           '-0', '-0.0e-999', '0e999999999', '1e999999999999', '1e-999999999999']
EDGES = ['4.9406564584124654e-324', '2.4703282292062327e-324', '2.4703282292062328e-324', '2.4703282292062329e-324',
         '2.2250738585072011e-308', '2.2250738585072014e-308', '1.7976931348623157e308', '1.7976931348623158e308',
         '1.7976931348623159e308', '1.797693134862315807e308', '1e309', '1e-400', '9007199254740993', '9007199254740992.5',
         '9007199254740993.0000000000000000000000001', '9007199254740992.99999999999999999999999', '0.1', '0.3', '100',
         '57.5', '1000000000000000000000', '123456789012345678901234567890', '0.000001', '1' + '0' * 400, '0.' + '0' * 400 + '1',
         '1' * 5000 + 'e-4900', '0.' + '0' * 3000 + '123456789012345678901234e3000']


def test_grammar_and_edges(checker):
    got = checker(GRAMMAR + EDGES)
    for tok, g in zip(GRAMMAR + EDGES, got):
        assert same(g, expected(tok)), (tok[:60], g, expected(tok))


def test_exact_ties_of_long_tokens(checker):
    """Halfway points between adjacent doubles written out exactly (up to ~770 digits), one digit
    above and a hair below: only the multi-word comparison can round these."""
    rng = random.Random(5)
    toks = []
    for _ in range(400):
        sub = rng.random() < 0.3
        m = rng.getrandbits(rng.randint(1, 52)) if sub else rng.getrandbits(53) | (1 << 52)
        e = -1074 if sub else rng.randint(-1074, 900)
        s = exact_decimal(Fraction(2 * m + 1) * Fraction(2) ** (e - 1))
        toks += [s, s + '1', s + '0' * 900 + '1']
        if s[-1] not in '.0':
            toks.append(s[:-1] + str(int(s[-1]) - 1) + '9' * 40)
    top = exact_decimal(Fraction(2 * ((1 << 53) - 1) + 1) * Fraction(2) ** 970)       # halfway to overflow
    toks += [top, top + '0001', top[:-2] + '1.' + '9' * 50]
    got = checker(toks)
    for tok, g in zip(toks, got):
        assert same(g, expected(tok)), (tok[:40], len(tok), g, expected(tok))


def test_random_tokens(checker):
    rng = random.Random(11)
    toks = []
    while len(toks) < 120000:
        r = rng.random()
        if r < .3:
            x = struct.unpack('<d', struct.pack('<Q', rng.getrandbits(64)))[0]
            if x == x and abs(x) != float('inf'):
                toks.append(repr(x))
        elif r < .5:
            toks.append('%d.%d' % (rng.randint(0, 100), rng.randint(0, 999999)))
        elif r < .7:
            toks.append(''.join(rng.choice('0123456789') for _ in range(rng.randint(1, 40))) + 'e%d' % rng.randint(-360, 330))
        elif r < .85:
            toks.append('0.' + ''.join(rng.choice('0123456789') for _ in range(rng.randint(17, 25))) + 'e%d' % rng.randint(-330, 310))
        else:
            x = struct.unpack('<d', struct.pack('<Q', rng.getrandbits(64) & 0x800FFFFFFFFFFFFF | rng.randint(0, 3) << 52))[0]
            toks.append('%.*e' % (rng.randint(15, 30), x))
    got = checker(toks)
    bad = [(t, g, expected(t)) for t, g in zip(toks, got) if not same(g, expected(t))]
    assert not bad, bad[:5]


def test_short_decimal_fast_path(checker):
    """parse_short_decimal: when it accepts a token completely it must give float()'s bits; it must
    accept the everyday forms (that is the point of it) and refuse what it cannot do exactly."""
    rng = random.Random(21)
    toks = ['0', '7', '100', '57.5', '-0', '-0.0', '5.', '.5', '-.5', '00012.50', '1700000000', '1700000000.123',
            '999999999999999', '9999999999999999', '0.000000000000001', '123456789.123456', '1234567890.123456',
            '.', '-', '-.', '', '1e5', '12a', '1_0', 'nan', '0.1', '0.3', '99.99999999999', '4.35', '2.675', '1.005']
    for _ in range(60000):
        a = rng.randint(0, 15)
        b = rng.randint(0, 15 - a)
        t = ''.join(rng.choice('0123456789') for _ in range(a))
        if b or rng.random() < .2:
            t += '.' + ''.join(rng.choice('0123456789') for _ in range(b))
        toks.append(('-' if rng.random() < .2 else '') + t)
    toks = [t for t in toks if '\n' not in t]
    checker(toks)
    accepted = 0
    for tok, (used, fbits) in zip(toks, checker.fast):
        if used == len(tok) and used:
            accepted += 1
            assert (1, fbits) == expected(tok), (tok, hex(fbits), expected(tok))
        else:
            digits = sum(ch.isdigit() for ch in tok)
            simple = digits and digits <= 15 and all(ch in '-.0123456789' for ch in tok) and tok.count('.') <= 1 \
                and '-' not in tok[1:]
            assert not simple, ('fast path refused a short decimal', tok, used)
    assert accepted > 50000
