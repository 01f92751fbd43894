"""The oracle is test infrastructure: nothing under spatial_queue_b200/ (the product) may import, load or execute
anything under oracle/, and bench.py may do so only in its CPU legs (reference arm / cpu_baseline)."""
import ast
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "spatial_queue_b200")


def _py_files(top):
    for d, _, files in os.walk(top):
        for f in files:
            if f.endswith(".py"):
                yield os.path.join(d, f)


def test_package_sources_never_name_the_oracle():
    bad = []
    for path in _py_files(PKG):
        tree = ast.parse(open(path).read(), path)
        for node in ast.walk(tree):
            names = []
            if isinstance(node, ast.Import):
                names = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                names = [node.module or ""]
            elif isinstance(node, ast.Call) and getattr(node.func, "id", getattr(node.func, "attr", "")) in ("import_module", "__import__"):
                names = [a.value for a in node.args if isinstance(a, ast.Constant) and isinstance(a.value, str)]
            bad += [(path, n) for n in names if n == "oracle" or n.startswith("oracle.")]
        # no path games either: the product never opens liboracle.so or oracle/_ref
        src = open(path).read()
        if re.search(r"liboracle|oracle/_ref|liblsp\.so", src):
            bad.append((path, "path"))
    for d, _, files in os.walk(os.path.join(PKG, "csrc")):
        for f in files:
            src = open(os.path.join(d, f)).read()
            if re.search(r'#include\s+"[^"]*oracle', src) or "liboracle" in src:
                bad.append((os.path.join(d, f), "include"))
    assert not bad, bad


def test_importing_the_package_does_not_load_the_oracle():
    code = ("import sys; sys.path.insert(0, %r); import spatial_queue_b200, spatial_queue_b200.sp_interface, "
            "spatial_queue_b200.tick, spatial_queue_b200.queue_class, spatial_queue_b200.synth, spatial_queue_b200.shard, "
            "spatial_queue_b200.flatten; "
            "bad = [m for m in sys.modules if m == 'oracle' or m.startswith('oracle.')]; "
            "print('BAD' if bad else 'CLEAN', bad)" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip().startswith("CLEAN"), out.stdout


def test_bench_touches_the_oracle_only_in_its_cpu_legs():
    """bench.py: every `oracle` import sits inside reference_arm, run_tick_cpu, the two reference helpers or the
    cpu_baseline block of main() -- never at module level and never in the timed GPU functions."""
    src = open(os.path.join(ROOT, "bench.py")).read()
    tree = ast.parse(src)
    allowed = {"_ref_worker", "sssp_reference_trees_per_s", "calibrate_ref_trees", "run_tick_cpu", "reference_arm", "main",
               "tick_reference_sample"}
    for node in tree.body:
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            mod = getattr(node, "module", None) or ",".join(a.name for a in node.names)
            assert "oracle" not in mod
    for fn in [n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef)]:
        uses = any(isinstance(n, ast.ImportFrom) and (n.module or "").startswith("oracle") for n in ast.walk(fn))
        if uses:
            assert fn.name in allowed, fn.name
    for name in ("run_tick_gpu", "make_tick_sim", "sssp_workload", "origins_for"):
        fn = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == name)
        assert "oracle" not in ast.get_source_segment(src, fn)
