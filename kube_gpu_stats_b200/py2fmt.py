"""Byte-compatible rendering of what the Python-2 reference prints.

* ``py2_repr``      - ``print line`` at gpu_util_stats.py:163: strings that came out of ``json``
                      are unicode under Py2, hence ``[u'node-1', u'pod-a', u'4', 37.5]``.
* ``py2_json_dump`` - ``json.dumps(d, indent=2)`` at gpu_util_stats.py:150: Py2.7 separates items
                      with ', ' so every line ending in a comma carries a trailing blank.
* ``pretty_table``  - prettytable's default style (who_use_gpu.py:11,21-22); the package is not
                      installed on this image and the reference does not pin a version.
"""
import json


def py2_repr(obj):
    if isinstance(obj, str):
        return 'u' + repr(obj)
    if isinstance(obj, list):
        return '[' + ', '.join(py2_repr(x) for x in obj) + ']'
    if isinstance(obj, tuple):
        inner = ', '.join(py2_repr(x) for x in obj)
        return '(' + inner + (',)' if len(obj) == 1 else ')')
    return repr(obj)


def py2_json_dump(obj):
    return json.dumps(obj, indent=2, separators=(', ', ': '))


def _centred(text, width):
    spare = width - len(text)
    left = spare // 2
    if spare % 2 and len(text) % 2 == 0:      # prettytable: odd leftover goes left for even-length text
        left += 1
    return ' ' * left + text + ' ' * (spare - left)


def pretty_table(header, rows):
    cells = [[str(v) for v in row] for row in rows]
    widths = [max([len(h)] + [len(r[i]) for r in cells]) for i, h in enumerate(header)]
    rule = '+' + '+'.join('-' * (w + 2) for w in widths) + '+'
    fmt = lambda vals: '|' + '|'.join(' ' + _centred(v, w) + ' ' for v, w in zip(vals, widths)) + '|'  # noqa: E731
    return '\n'.join([rule, fmt(header), rule] + [fmt(r) for r in cells] + [rule])
