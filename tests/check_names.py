"""Developer tool: a coarse undefined-name check (scope-insensitive: a name that is loaded somewhere must be bound
somewhere in the same file or be a builtin) for the scripts that only ever run on a GPU box - a typo there costs a GPU
session. No third-party analyzer is installed in this image.
    python tests/check_names.py [files...]      default: the device-only scripts and timing tools
"""
import ast
import builtins
import glob
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bound_and_loaded(tree):
    bound, loaded = set(dir(builtins)) | {"__file__", "__name__"}, []
    for node in ast.walk(tree):
        if isinstance(node, ast.Name):
            if isinstance(node.ctx, ast.Load):
                loaded.append((node.id, node.lineno))
            else:
                bound.add(node.id)
        elif isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            bound.add(node.name)
            if not isinstance(node, ast.ClassDef):
                a = node.args
                for arg in a.posonlyargs + a.args + a.kwonlyargs + ([a.vararg] if a.vararg else []) + ([a.kwarg] if a.kwarg else []):
                    bound.add(arg.arg)
        elif isinstance(node, ast.Lambda):
            a = node.args
            for arg in a.posonlyargs + a.args + a.kwonlyargs + ([a.vararg] if a.vararg else []) + ([a.kwarg] if a.kwarg else []):
                bound.add(arg.arg)
        elif isinstance(node, (ast.Import, ast.ImportFrom)):
            for alias in node.names:
                bound.add((alias.asname or alias.name).split(".")[0])
        elif isinstance(node, ast.ExceptHandler) and node.name:
            bound.add(node.name)
        elif isinstance(node, (ast.Global, ast.Nonlocal)):
            bound.update(node.names)
    return bound, loaded


def check(path):
    tree = ast.parse(open(path).read(), path)
    bound, loaded = bound_and_loaded(tree)
    return [(name, line) for name, line in loaded if name not in bound]


def default_files():
    pats = ["tests/dist_*_check.py", "tests/wide_sgemm_probe.py", "tests/test_zz_*.py", "tests/gpu_util.py",
            "profiles/tools/*.py", "bench.py", "__graft_entry__.py"]
    return sorted(f for p in pats for f in glob.glob(os.path.join(ROOT, p)))


def main():
    files = sys.argv[1:] or default_files()
    bad = 0
    for f in files:
        for name, line in check(f):
            print(f"{os.path.relpath(f, ROOT)}:{line}: name '{name}' is never bound in this file")
            bad += 1
    print(f"checked {len(files)} files, {bad} unbound names")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
