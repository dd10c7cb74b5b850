"""Oracle restatement of the reference's ``who_use_gpu/who_use_gpu.py``.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  ``WUG:a-b`` =
``/root/reference/who_use_gpu/who_use_gpu.py`` lines a..b.  The ``kubectl``
subprocess (WUG:8-10) is replaced by the decoded JSON it would have produced.
**Parity unpinned**: the reference has no tests and is Python 2 only; the rules
are frozen by ``tests/golden/ref_rules_wug.json``.  ``prettytable`` is a
third-party dependency of the reference (WUG:5, version unpinned, not installed
here); ``render_table`` restates its documented default style.
"""

GPU_LIMIT_KEY = 'alpha.kubernetes.io/nvidia-gpu'      # WUG:40 (legacy key, not nvidia.com/gpu)
HEADER = ['Namespace', 'Node', 'Pod', 'GPU Type', 'GPU Cores']   # WUG:11


def gpu_type_of(pod):
    """WUG:51-58: first value of the first required node-affinity expression."""
    try:
        return (pod['spec']['affinity']['nodeAffinity']
                ['requiredDuringSchedulingIgnoredDuringExecution']
                ['nodeSelectorTerms'][0]['matchExpressions'][0]['values'][0])
    except Exception:
        return '<unspecified>'


def identity_record(pod, printed):
    """WUG:27-49 ``checkPod``.  ``printed`` collects what the reference prints when
    it swallows an exception (WUG:46-48): ``print(e)`` of a KeyError shows the
    quoted key, e.g. ``'nodeName'``."""
    try:
        meta = pod['metadata']
        name = meta['name']
        namespace = meta['namespace']
        containers = pod['spec']['containers']
        node = pod['spec']['nodeName']            # read before GPUs are counted (WUG:33)
        gpus = 0
        for c in containers:
            if 'resources' not in c or 'limits' not in c['resources']:
                continue
            gpus += int(c['resources']['limits'].get(GPU_LIMIT_KEY, 0))
        if gpus > 0:
            return [namespace, node, name, gpu_type_of(pod), gpus]
    except Exception as exc:
        printed.append(str(exc))
        return []
    return []


def ownership(pods_json):
    """WUG:12-21: rows, grand total and per-type totals over *emitted* rows."""
    printed, rows, per_type, total = [], [], {}, 0
    for pod in pods_json['items']:
        row = identity_record(pod, printed)
        if not row:
            continue
        total += row[4]
        per_type[row[3]] = per_type.get(row[3], 0) + row[4]
        rows.append(row)
    return rows, total, per_type, printed


def _centre(text, width):
    # prettytable's centring: odd leftover goes right for odd-length text, left for even.
    spare = width - len(text)
    left = spare // 2
    if spare % 2 and len(text) % 2 == 0:
        left += 1
    return ' ' * left + text + ' ' * (spare - left)


def render_table(header, rows):
    """prettytable default style (WUG:22): '+---+' rules, '|' bars, one blank of
    padding, every cell centred.  An empty table still prints header and frame."""
    cells = [[str(v) for v in r] for r in rows]
    widths = [max([len(h)] + [len(r[i]) for r in cells]) for i, h in enumerate(header)]
    rule = '+' + '+'.join('-' * (w + 2) for w in widths) + '+'

    def line(vals):
        return '|' + '|'.join(' ' + _centre(v, w) + ' ' for v, w in zip(vals, widths)) + '|'

    return '\n'.join([rule, line(header), rule] + [line(r) for r in cells] + [rule])


def main_text(pods_json):
    """WUG:7-25 as text.  Per-type lines follow first-seen order (Py2 dict order
    is unspecified, WUG:24)."""
    rows, total, per_type, printed = ownership(pods_json)
    out = list(printed)
    out.append(render_table(HEADER, rows))
    out.append('Total GPU: %d' % total)
    out.extend('%s\t%d' % kv for kv in per_type.items())
    return '\n'.join(out) + '\n'
