"""Every Python file of the repository parses (the GPU check scripts are only imported on a GPU box, so a syntax error in them
would otherwise show up there first)."""
import ast
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_all_python_files_parse():
    bad = []
    for top in ("", "scripts", "examples", "tests", "cori_intml_examples_b200"):
        base = os.path.join(ROOT, top)
        for dirpath, dirs, files in os.walk(base):
            dirs[:] = [d for d in dirs if d not in ("__pycache__", "gpurun_out", "baseline", ".git")] if top else []
            for f in files:
                if f.endswith(".py"):
                    path = os.path.join(dirpath, f)
                    try:
                        ast.parse(open(path, encoding="utf8").read(), path)
                    except SyntaxError as e:
                        bad.append("%s: %s" % (path, e))
    assert not bad, bad
