"""Graph visualisation: render ``executor.topo_order`` to Graphviz DOT (+ PNG/SVG when the ``dot``
binary or the graphviz python package is available) and optionally serve the directory over HTTP.

reference: python/graphboard/graph2fig.py:11-32 (graphviz Digraph of the topo order, PNG, then
``python -m SimpleHTTPServer``).  File rendering only - no GL / display involved.
"""
from __future__ import annotations

import os
import shutil
import subprocess

_COLORS = {'PlaceholderOp': '#cfe8ff', 'OptimizerOp': '#ffd9b3', 'DataloaderOp': '#d9f2d9'}


def to_dot(executor, name='hetu_graph') -> str:
    nodes = executor.topo_order if hasattr(executor, 'topo_order') else list(executor)
    ids = {n: 'n%d' % i for i, n in enumerate(nodes)}
    lines = ['digraph "%s" {' % name, '  rankdir=TB;', '  node [shape=box, style="rounded,filled", fontname="Helvetica"];']
    for n in nodes:
        color = _COLORS.get(type(n).__name__, _COLORS.get(n.op_type, '#f2f2f2'))
        if 'Communicate' in n.op_type or 'Pipeline' in n.op_type or n.op_type.startswith('MP'):
            color = '#f9c2c2'
        label = n.name + ('\\n%s' % (n.ctx,) if getattr(n, 'ctx', None) is not None else '')
        lines.append('  %s [label="%s", fillcolor="%s"];' % (ids[n], label.replace('"', "'"), color))
    for n in nodes:
        for i in n.inputs:
            if i is not None and i in ids:
                lines.append('  %s -> %s;' % (ids[i], ids[n]))
    lines.append('}')
    return '\n'.join(lines)


def show(executor, path='graphboard_out', fmt='png', serve=False, port=9997):
    """Write ``<path>/output.dot`` (and ``output.<fmt>`` when Graphviz is installed); with
    ``serve=True`` start ``http.server`` on ``port`` like the reference does."""
    os.makedirs(path, exist_ok=True)
    dot_file = os.path.join(path, 'output.dot')
    with open(dot_file, 'w') as f:
        f.write(to_dot(executor))
    rendered = None
    if shutil.which('dot'):
        rendered = os.path.join(path, 'output.' + fmt)
        subprocess.run(['dot', '-T' + fmt, dot_file, '-o', rendered], check=False)
    else:
        try:
            import graphviz                                            # noqa: F401
            rendered = graphviz.Source(open(dot_file).read()).render(os.path.join(path, 'output'), format=fmt)
        except Exception:
            rendered = None
    with open(os.path.join(path, 'index.html'), 'w') as f:
        f.write('<html><body><h3>hetu graph</h3>%s<pre>%s</pre></body></html>' % (
            '<img src="output.%s"/>' % fmt if rendered else '<p>(Graphviz not installed: DOT source below)</p>',
            open(dot_file).read().replace('<', '&lt;')))
    if serve:
        subprocess.Popen(['python', '-m', 'http.server', str(port)], cwd=path)
    return rendered or dot_file
