"""Names used but never bound anywhere in their module (GPU-only branches are not executed by the CPU suite, so a
missing import there would otherwise first show up on the GPU box)."""
import ast
import builtins
import pathlib

ROOT = pathlib.Path(__file__).resolve().parents[1]


def _unbound(path: pathlib.Path):
    tree = ast.parse(path.read_text())
    bound = set(dir(builtins)) | {"__file__", "__name__", "__doc__"}
    for node in ast.walk(tree):
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            bound.update((a.asname or a.name).split(".")[0] for a in node.names)
        elif isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            bound.add(node.name)
        elif isinstance(node, ast.arg):
            bound.add(node.arg)
        elif isinstance(node, ast.Name) and isinstance(node.ctx, (ast.Store, ast.Del)):
            bound.add(node.id)
        elif isinstance(node, ast.ExceptHandler) and node.name:
            bound.add(node.name)
    return [(node.lineno, node.id) for node in ast.walk(tree)
            if isinstance(node, ast.Name) and isinstance(node.ctx, ast.Load) and node.id not in bound]


def test_no_unbound_names_in_the_package_and_benches():
    files = list((ROOT / "tiresias_b200").rglob("*.py")) + list((ROOT / "bench").glob("*.py")) + \
        [ROOT / "bench.py", ROOT / "__graft_entry__.py"]
    bad = {str(f.relative_to(ROOT)): _unbound(f) for f in files}
    bad = {k: v for k, v in bad.items() if v}
    assert not bad, bad
