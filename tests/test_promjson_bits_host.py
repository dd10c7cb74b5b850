"""The chunk lexer of csrc/promjson_bits.cuh (bit arithmetic on 64-byte chunks of a Prometheus
answer) compiled as plain C++ and checked against a byte-by-byte loop: escapes, strings, bracket
depth under both start states, and the structural events.  CPU-only check of the arithmetic."""
import random
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope='module')
def lexer(tmp_path_factory):
    exe = tmp_path_factory.mktemp('bits') / 'promjson_bits_check'
    subprocess.run(['g++', '-O2', '-std=c++17', '-o', str(exe), str(ROOT / 'tests/host/promjson_bits_check.cpp')], check=True)

    def run(cases):
        text = ''.join('%s %d %d %d %d\n' % (c[0].hex(), c[1], c[2], c[3], c[4]) for c in cases)
        out = subprocess.run([str(exe)], input=text, capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    return run


def by_bytes(chunk, first_escaped, in_string, depth, pair_depth):
    """The specification: one byte at a time."""
    esc_mask = inside = opened = closed = 0
    esc, par, d = bool(first_escaped), 0, [0, 0]
    events = []
    ins = in_string
    for i, c in enumerate(chunk):
        if esc:
            esc_mask |= 1 << i
            esc = False
            if par:
                inside |= 1 << i
            continue
        if c == 0x5C:
            esc = True
            if par:
                inside |= 1 << i
            continue
        if c == 0x22:
            par ^= 1
            ins ^= 1
        elif c in b'[{':
            opened |= 1 << i
            d[par] += 1
            if not ins:
                if c == 0x5B:
                    if depth == pair_depth:
                        events.append((0, i))
                    elif depth == 4:
                        events.append((2, i))
                elif depth == 3:
                    events.append((1, i))
                depth += 1
        elif c in b']}':
            closed |= 1 << i
            d[par] -= 1
            if not ins:
                depth -= 1
                if c == 0x5D:
                    if depth == 4 and pair_depth == 5:
                        events.append((3, i))
                elif depth == 3:
                    events.append((4, i))
        if par:
            inside |= 1 << i
    return esc_mask, inside, opened, closed, par, d[0], d[1], events


def test_chunk_lexer_equals_byte_loop(lexer):
    rng = random.Random(7)
    alphabets = [b'\\"[]{}a,:1 ', b'\\\\\\"a', b'[]{}"', bytes(range(256)), b'[1700000000,"57.5"],', b'\\']
    cases = []
    for _ in range(6000):
        alpha = rng.choice(alphabets)
        chunk = bytes(rng.choice(alpha) for _ in range(64))
        cases.append((chunk, rng.randint(0, 1), rng.randint(0, 1), rng.randint(0, 7), rng.choice((4, 5))))
    sample = b'{"metric":{"pod_name":"a\\"b"},"values":[[1700000000,"57.5"],[1700000001,"3"]]},{"metric":{}'
    for shift in range(0, len(sample) - 64):
        cases.append((sample[shift:shift + 64], 0, 0, 3, 5))
    got = lexer(cases)
    assert len(got) == len(cases)
    for case, line in zip(cases, got):
        esc, inside, opened, closed, flip, d0, d1, events = by_bytes(*case)
        head, _, ev = line.partition('|')
        f = head.split()
        g_events = [tuple(int(x) for x in e.split(':')) for e in ev.split()]
        assert int(f[0], 16) == esc, ('escaped', case)
        assert int(f[1], 16) == inside, ('inside', case)
        assert int(f[2], 16) == opened and int(f[3], 16) == closed, ('brackets', case)
        assert (int(f[4]), int(f[5]), int(f[6])) == (flip, d0, d1), ('transition', case)
        assert g_events == events, ('events', case, g_events, events)
