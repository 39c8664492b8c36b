"""Case tables shared by tests and tests/golden/make_golden.py (kept in one place there)."""
import importlib.util
import os

_spec = importlib.util.spec_from_file_location(
    '_make_golden_tables', os.path.join(os.path.dirname(__file__), 'golden', 'make_golden.py'))


def _tables():
    # parse only the two literal tables; importing make_golden would import the reference
    import ast
    src = open(_spec.origin).read()
    tree = ast.parse(src)
    out = {}
    for node in tree.body:
        if isinstance(node, ast.Assign) and node.targets[0].id in ('STFT_CASES', 'CONV_CASES'):
            out[node.targets[0].id] = eval(compile(ast.Expression(node.value), 'tables', 'eval'))
    return out


_T = _tables()
STFT_CASES = _T['STFT_CASES']
CONV_CASES = _T['CONV_CASES']
