"""Undefined-name check without third-party linters: every name a function or class body reads from the module scope must be
assigned, imported or defined somewhere in that module (or be a builtin).  Catches typos in branches the tests do not reach.

    python tools/check_names.py [paths...]        # default: fedml_b200 fedml tools bench.py __graft_entry__.py
"""
import builtins
import os
import symtable
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def module_names(table):
    return {s.get_name() for s in table.get_symbols() if s.is_assigned() or s.is_imported() or s.is_namespace()}


def walk(table, defined, path, problems):
    for child in table.get_children():
        for s in child.get_symbols():
            n = s.get_name()
            if s.is_referenced() and s.is_global() and not s.is_assigned() and n not in defined and not hasattr(builtins, n) \
                    and n not in ("__file__", "__name__", "__doc__", "__class__", "__path__", "__spec__", "__package__", "__builtins__"):
                problems.append(f"{path}: {child.get_name()} (line {child.get_lineno()}) reads undefined name {n!r}")
        walk(child, defined, path, problems)


def check(path, problems):
    with open(path, encoding="utf-8") as fh:
        src = fh.read()
    try:
        table = symtable.symtable(src, path, "exec")
    except SyntaxError as exc:
        problems.append(f"{path}: syntax error: {exc}")
        return
    if "import *" in src:
        return
    walk(table, module_names(table), os.path.relpath(path, ROOT), problems)


def main():
    targets = sys.argv[1:] or [os.path.join(ROOT, p) for p in ("fedml_b200", "fedml", "tools", "baseline", "examples", "tests", "bench.py", "__graft_entry__.py")]
    problems = []
    for t in targets:
        if os.path.isfile(t):
            check(t, problems)
            continue
        for dp, dn, fn in os.walk(t):
            dn[:] = [d for d in dn if d not in ("__pycache__", "_ref")]
            for f in fn:
                if f.endswith(".py"):
                    check(os.path.join(dp, f), problems)
    print("\n".join(problems) or "no undefined names")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main())
