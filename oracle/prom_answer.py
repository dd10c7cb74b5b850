"""ORACLE (test infrastructure only - never imported by kube_gpu_stats_b200).

Plain-Python statement of what ``kgs_matrix_parse`` reports about the TEXT of a Prometheus answer
(the numbers themselves are ``float()``, gpu_util_stats.py:48,:74).  One pass, one byte at a time:

* a character behind an odd run of backslashes is escaped; an unescaped ``"`` toggles "inside a string";
* outside strings ``{`` / ``[`` open a level and ``}`` / ``]`` close one;
* a ``{`` that opens at depth 3 starts a series (``data.result[i]``), the ``[`` that opens at depth 4
  inside it is its ``values`` array, a ``[`` that opens at depth 5 is a ``[ts, "v"]`` pair
  (for an instant ``vector`` the pair itself opens at depth 4).

Returns the byte position of every pair, the CSR offsets (first pair of every series) and per series
the four byte spans ``{``, ``[`` of values, ``]`` of values, ``}`` (2**64-1 where there is none)."""
NONE = 0xFFFFFFFFFFFFFFFF


def scan_answer(raw, vector=False):
    pair_depth = 4 if vector else 5
    pos, offsets, spans = [], [], []
    esc = in_string = False
    depth = 0
    for i, c in enumerate(raw):
        if esc:
            esc = False
            continue
        if c == 0x5C:
            esc = True
            continue
        if c == 0x22:
            in_string = not in_string
            continue
        if in_string:
            continue
        if c == 0x5B:                                   # [
            if depth == pair_depth:
                pos.append(i)
            elif depth == 4 and spans:
                spans[-1][1] = i
            depth += 1
        elif c == 0x7B:                                 # {
            if depth == 3:
                offsets.append(len(pos))
                spans.append([i, NONE, NONE, NONE])
            depth += 1
        elif c == 0x5D:                                 # ]
            depth -= 1
            if depth == 4 and not vector and spans:
                spans[-1][2] = i
        elif c == 0x7D:                                 # }
            depth -= 1
            if depth == 3 and spans:
                spans[-1][3] = i
    if in_string or depth != 0:
        raise ValueError('the text ends inside a string or with unbalanced brackets')
    offsets.append(len(pos))
    return pos, offsets, spans
