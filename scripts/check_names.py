#!/usr/bin/env python
"""Scope-aware undefined-name check over the Python sources (no third-party linter in the image).

    python scripts/check_names.py [paths...]

Most of the device-side Python only runs on a GPU box; a misspelt or un-imported name there would first show up
in the round-end GPU run. This walks every module with `ast`, keeps the chain of enclosing scopes (module,
functions, lambdas, comprehensions, class bodies) and reports every name that is read but bound nowhere visible.
A second pass resolves every alias of one of this repository's own modules (`from ..ops import _cuda as K`,
also inside functions) and checks that `K.<attr>` exists in that module (all of them import without a GPU).
A third pass binds the arguments of every call of one of the repository's own functions or classes
(`coll.all_reduce(comm, buf, n, ..., out=stage)`) against its signature: wrong arity or an unknown keyword;
a fourth does the same for `self.method(...)` inside the package's classes.
Exit code 1 if anything is found."""
import ast
import builtins
import os
import sys

def bound_names(node):
    """Names bound directly in this scope (nested function / class / lambda / comprehension scopes are not entered)."""
    out = set()

    def visit(n):
        for ch in ast.iter_child_nodes(n):
            if isinstance(ch, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
                out.add(ch.name)
                for d in ch.decorator_list:
                    visit(d)
                continue
            if isinstance(ch, (ast.Lambda, ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
                continue
            if isinstance(ch, ast.Import):
                out.update((a.asname or a.name).split('.')[0] for a in ch.names)
            elif isinstance(ch, ast.ImportFrom):
                out.update(a.asname or a.name for a in ch.names)
            elif isinstance(ch, ast.Name) and isinstance(ch.ctx, (ast.Store, ast.Del)):
                out.add(ch.id)
            elif isinstance(ch, ast.ExceptHandler) and ch.name:
                out.add(ch.name)
            elif isinstance(ch, (ast.Global, ast.Nonlocal)):
                out.update(ch.names)
            visit(ch)

    visit(node)
    return out


def args_of(fn):
    a = fn.args
    out = {x.arg for x in a.args + a.kwonlyargs + a.posonlyargs}
    for extra in (a.vararg, a.kwarg):
        if extra:
            out.add(extra.arg)
    return out


bad = 0


def check(node, scopes, path):
    """Every name read under `node` must be bound in one of `scopes` (innermost last)."""
    global bad
    for ch in ast.iter_child_nodes(node):
        if isinstance(ch, (ast.FunctionDef, ast.AsyncFunctionDef)):
            for d in ch.decorator_list + ch.args.defaults + [x for x in ch.args.kw_defaults if x]:
                check_expr(d, scopes, path)
            check(ch, scopes + [bound_names(ch) | args_of(ch)], path)
        elif isinstance(ch, ast.Lambda):
            check(ch, scopes + [args_of(ch)], path)
        elif isinstance(ch, ast.ClassDef):
            # names of a class body are visible in the body itself only; its methods see the enclosing scopes
            body_names = bound_names(ch)
            for b in ch.bases + ch.decorator_list:
                check_expr(b, scopes, path)
            for stmt in ch.body:
                inner = scopes if isinstance(stmt, (ast.FunctionDef, ast.AsyncFunctionDef)) else scopes + [body_names]
                check(ast.Module(body=[stmt], type_ignores=[]), inner, path)
        elif isinstance(ch, (ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
            names = {n.id for g in ch.generators for n in ast.walk(g.target) if isinstance(n, ast.Name)}
            check(ch, scopes + [names], path)
        elif isinstance(ch, ast.Name) and isinstance(ch.ctx, ast.Load):
            if not any(ch.id in s for s in scopes):
                print('%s:%d undefined name %s' % (path, ch.lineno, ch.id))
                bad += 1
        else:
            check(ch, scopes, path)


def check_expr(e, scopes, path):
    check(ast.Module(body=[ast.Expr(e)], type_ignores=[]), scopes, path)


def python_files(base):
    return [base] if base.endswith('.py') else [os.path.join(r, f) for r, _, fs in os.walk(base) for f in fs if f.endswith('.py')]


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT = ['kungfu_b200', 'kungfu', 'tests', 'benchmarks', 'examples', 'scripts', 'bench.py', '__graft_entry__.py', 'setup.py']
TARGETS = sys.argv[1:] or [os.path.join(ROOT, d) for d in DEFAULT]
MODULE_GLOBALS = set(dir(builtins)) | {'__file__', '__name__', '__doc__', '__builtins__', '__spec__', '__path__'}
for base in TARGETS:
    for p in python_files(base):
        with open(p) as fh:
            tree = ast.parse(fh.read())
        check(tree, [bound_names(tree) | MODULE_GLOBALS], p)
print('undefined names: %d' % bad)


# ---- second pass: attributes of this repository's modules ------------------------------------------
import importlib  # noqa: E402

sys.path.insert(0, ROOT)
missing = 0


def _module_name(path):
    rel = os.path.relpath(path, ROOT)[:-3].replace(os.sep, '.')
    return rel[:-9] if rel.endswith('.__init__') else rel


def _resolve(frm, level, pkg):
    if level == 0:
        return frm
    parts = pkg.split('.')
    return '.'.join(parts[:len(parts) - (level - 1)] + ([frm] if frm else []))


for base in TARGETS:
    for path in python_files(base):
        tree = ast.parse(open(path).read())
        mod = _module_name(path)
        pkg = mod if path.endswith('__init__.py') else (mod.rsplit('.', 1)[0] if '.' in mod else '')
        aliases = {}
        for node in ast.walk(tree):
            if isinstance(node, ast.Import):
                for a in node.names:
                    if a.name.startswith('kungfu_b200') and a.asname:
                        aliases[a.asname] = a.name
            elif isinstance(node, ast.ImportFrom):
                src = _resolve(node.module or '', node.level, pkg) if node.level else (node.module or '')
                if not src.startswith('kungfu_b200'):
                    continue
                for a in node.names:
                    try:
                        importlib.import_module(src + '.' + a.name)  # only names that ARE modules
                        aliases[a.asname or a.name] = src + '.' + a.name
                    except Exception:
                        pass
        mods = {k: importlib.import_module(v) for k, v in aliases.items()}
        rebound = {n.id for n in ast.walk(tree) if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Store)}
        for node in ast.walk(tree):
            if isinstance(node, ast.Attribute) and isinstance(node.value, ast.Name) and isinstance(node.ctx, ast.Load) \
                    and node.value.id in mods and node.value.id not in rebound and not hasattr(mods[node.value.id], node.attr):
                print('%s:%d %s.%s: module %s has no such attribute' %
                      (path, node.lineno, node.value.id, node.attr, mods[node.value.id].__name__))
                missing += 1
print('missing module attributes: %d' % missing)

# ---- third pass: calls of this repository's own functions / classes against their signatures ----------
import inspect  # noqa: E402

checked = mismatched = 0
for base in TARGETS:
    for path in python_files(base):
        tree = ast.parse(open(path).read())
        mod = _module_name(path)
        pkg = mod if path.endswith('__init__.py') else (mod.rsplit('.', 1)[0] if '.' in mod else '')
        objs = {}
        for node in ast.walk(tree):
            if isinstance(node, ast.Import):
                for a in node.names:
                    if a.name.startswith('kungfu_b200') and a.asname:
                        objs[a.asname] = importlib.import_module(a.name)
            elif isinstance(node, ast.ImportFrom):
                src = _resolve(node.module or '', node.level, pkg) if node.level else (node.module or '')
                if not src.startswith('kungfu_b200'):
                    continue
                try:
                    m = importlib.import_module(src)
                except Exception:
                    continue
                for a in node.names:
                    try:
                        objs[a.asname or a.name] = importlib.import_module(src + '.' + a.name)
                    except Exception:
                        if hasattr(m, a.name):
                            objs[a.asname or a.name] = getattr(m, a.name)
        rebound = {n.id for n in ast.walk(tree) if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Store)}
        rebound |= {n.name for n in ast.walk(tree) if isinstance(n, (ast.FunctionDef, ast.ClassDef))}
        for node in ast.walk(tree):
            if not isinstance(node, ast.Call) or any(isinstance(a, ast.Starred) for a in node.args) \
                    or any(k.arg is None for k in node.keywords):
                continue
            f, target = node.func, None
            if isinstance(f, ast.Attribute) and isinstance(f.value, ast.Name) and f.value.id in objs and f.value.id not in rebound:
                o = objs[f.value.id]
                if inspect.ismodule(o) and hasattr(o, f.attr):
                    target = getattr(o, f.attr)
            elif isinstance(f, ast.Name) and f.id in objs and f.id not in rebound:
                target = objs[f.id]
            if inspect.isclass(target) and inspect.isfunction(target.__init__):
                sig, args = inspect.signature(target.__init__), [None] * (len(node.args) + 1)
            elif inspect.isfunction(target):
                sig, args = inspect.signature(target), [None] * len(node.args)
            else:
                continue
            checked += 1
            try:
                sig.bind(*args, **{k.arg: None for k in node.keywords})
            except TypeError as e:
                print('%s:%d %s(...): %s' % (path, node.lineno, ast.unparse(f), e))
                mismatched += 1
print('calls of own functions checked: %d, not matching the signature: %d' % (checked, mismatched))

# ---- fourth pass: self.method(...) inside the package's classes against the method found on the class ----
# (methods that a dynamically composed base provides -- the optimizer mix-ins -- are not resolvable here and are skipped)
self_checked = self_bad = 0
pkg_root = os.path.join(ROOT, 'kungfu_b200')
for r, _, fs in os.walk(pkg_root):
    for fname in fs:
        if not fname.endswith('.py') or fname == '__main__.py':
            continue
        path = os.path.join(r, fname)
        try:
            m = importlib.import_module(_module_name(path))
        except Exception:
            continue
        for cls in [n for n in ast.parse(open(path).read()).body if isinstance(n, ast.ClassDef)]:
            c = getattr(m, cls.name, None)
            if not inspect.isclass(c):
                continue
            for node in ast.walk(cls):
                if not (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and isinstance(node.func.value, ast.Name)
                        and node.func.value.id == 'self' and hasattr(c, node.func.attr)):
                    continue
                if any(isinstance(a, ast.Starred) for a in node.args) or any(k.arg is None for k in node.keywords):
                    continue
                t = inspect.getattr_static(c, node.func.attr)
                fn = t.__func__ if isinstance(t, (staticmethod, classmethod)) else t
                if not inspect.isfunction(fn):
                    continue
                self_checked += 1
                try:
                    inspect.signature(fn).bind(*([None] * (len(node.args) + (0 if isinstance(t, staticmethod) else 1))),
                                               **{k.arg: None for k in node.keywords})
                except TypeError as e:
                    print('%s:%d self.%s(...): %s' % (path, node.lineno, node.func.attr, e))
                    self_bad += 1
print('self.method calls checked: %d, not matching: %d' % (self_checked, self_bad))
sys.exit(1 if bad or missing or mismatched or self_bad else 0)
