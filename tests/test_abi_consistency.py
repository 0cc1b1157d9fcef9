"""Python <-> C ABI consistency, checked without a GPU.

The ctypes tables (kungfu_b200/ops/_cuda.py:_SIGS, kungfu_b200/_host.py:_declare) restate the prototypes of
include/kungfu_b200.h and include/kungfu_host.h by hand, and the device-side call sites (`K.call('kfb_...', ...)`)
restate them once more. A drift between any two of them is undefined behaviour that only a GPU run would show
(an int passed where the kernel entry expects an int64, a missing stream argument). This parses the headers and the
Python sources and compares: every declared entry point exists in the header and in the built library, with the same
number of parameters of the same class (pointer / 32-bit / 64-bit / float), and every call site passes that many.
"""
import ast
import ctypes as C
import os
import re

from conftest import ROOT


def _split_params(s):
    out, depth, cur = [], 0, ''
    for ch in s:
        if ch == '(':
            depth += 1
        elif ch == ')':
            depth -= 1
        if ch == ',' and depth == 0:
            out.append(cur.strip())
            cur = ''
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return [] if out == ['void'] else out


def _c_class(decl):
    """pointer / i32 / i64 / f32 / f64 of one C parameter or return declaration."""
    d = re.sub(r'\bconst\b|\bstruct\b', ' ', decl)
    if '(*' in d or '*' in d or '[' in d:
        return 'ptr'
    d = re.sub(r'\b[a-zA-Z_]\w*\s*$', '', d.strip()) if len(d.split()) > 1 else d  # drop the parameter name
    d = ' '.join(d.split())
    if d in ('cudaStream_t', 'kfb_stream_t'):
        return 'ptr'
    if d in ('int', 'int32_t', 'unsigned', 'unsigned int', 'uint32_t'):
        return 'i32'
    if d in ('int64_t', 'uint64_t', 'long long', 'unsigned long long', 'size_t', 'long', 'unsigned long'):
        return 'i64'
    if d == 'float':
        return 'f32'
    if d == 'double':
        return 'f64'
    if d == 'void':
        return 'void'
    raise AssertionError('unclassified C type %r (from %r)' % (d, decl))


def _py_class(t):
    if t is None:
        return 'void'
    if t in (C.c_void_p, C.c_char_p) or hasattr(t, 'contents') or issubclass(t, C._Pointer):
        return 'ptr'
    if t in (C.c_int, C.c_int32, C.c_uint, C.c_uint32):
        return 'i32'
    if t in (C.c_int64, C.c_longlong, C.c_ulonglong, C.c_size_t, C.c_uint64, C.c_long, C.c_ulong):
        return 'i64'
    if t is C.c_float:
        return 'f32'
    if t is C.c_double:
        return 'f64'
    raise AssertionError('unclassified ctypes type %r' % (t, ))


def _header_prototypes(path, prefix):
    text = open(path).read()
    text = re.sub(r'/\*.*?\*/', ' ', text, flags=re.S)
    text = re.sub(r'//[^\n]*', ' ', text)
    text = re.sub(r'^\s*#[^\n]*(\\\n[^\n]*)*', ' ', text, flags=re.M)  # preprocessor lines
    text = re.sub(r'extern\s+"C"\s*\{|^\s*\}\s*$', ' ', text, flags=re.M)
    protos = {}
    for m in re.finditer(r'([A-Za-z_][\w\s\*]*?)\b(%s\w+)\s*\(([^;{]*)\)\s*;' % prefix, text):
        ret, name, params = m.group(1).strip(), m.group(2), m.group(3)
        if ret.startswith('typedef') or 'return' in ret:
            continue
        protos[name] = (ret, _split_params(' '.join(params.split())))
    return protos


def _check_table(table, protos, lib, where):
    problems = []
    for name, (res, args) in sorted(table.items()):
        if name not in protos:
            problems.append('%s: %s is declared in Python but not in the header' % (where, name))
            continue
        if not hasattr(lib, name):
            problems.append('%s: %s is not exported by the built library' % (where, name))
        ret, params = protos[name]
        if len(params) != len(args):
            problems.append('%s: %s takes %d parameters in the header, %d in Python' % (where, name, len(params), len(args)))
            continue
        for i, (cp, pt) in enumerate(zip(params, args)):
            if _c_class(cp) != _py_class(pt):
                problems.append('%s: %s parameter %d is `%s` in the header, %s in Python' % (where, name, i, cp, pt.__name__))
        rc, rp = _c_class(ret + ' x') if ret != 'void' else 'void', _py_class(res)
        if ('*' in ret and rp != 'ptr') or ('*' not in ret and rc != rp):
            problems.append('%s: %s returns `%s` in the header, %s in Python' % (where, name, ret, getattr(res, '__name__', res)))
    return problems


def test_device_ctypes_table_matches_the_header_and_the_library():
    from kungfu_b200.ops import _cuda as K
    protos = _header_prototypes(os.path.join(ROOT, 'include', 'kungfu_b200.h'), 'kfb_')
    assert len(protos) > 60, len(protos)
    problems = _check_table(K._SIGS, protos, K.lib(), '_cuda.py')
    undeclared = sorted(n for n in protos if n not in K._SIGS and n not in ('kfb_fuse_noalloc', ))
    assert not problems, '\n'.join(problems)
    # every entry point of the header that Python may call by name has a signature (the no-alloc gather is C-only)
    assert not undeclared, 'in the header but without a ctypes signature: %s' % undeclared


def test_host_ctypes_table_matches_the_header_and_the_library():
    from kungfu_b200 import _host as H
    protos = _header_prototypes(os.path.join(ROOT, 'include', 'kungfu_host.h'), 'kf_')
    assert len(protos) > 50, len(protos)
    lib = H.lib()
    src = open(os.path.join(ROOT, 'kungfu_b200', '_host.py')).read()
    names = set(re.findall(r"\bf\('(kf_\w+)'", src))
    for tup in re.findall(r"for n in \(([^)]*)\):\s*\n\s*f\(n,", src):
        names |= set(re.findall(r"'(kf_\w+)'", tup))
    assert len(names) > 40, len(names)
    table = {n: (getattr(lib, n).restype, list(getattr(lib, n).argtypes or [])) for n in names}
    problems = _check_table(table, protos, lib, '_host.py')
    assert not problems, '\n'.join(problems)


def _call_sites():
    """(file, line, entry point, number of arguments) of every K.call('kfb_x', ...) / <lib>.kfb_x(...) / <lib>.kf_x(...)."""
    out = []
    for base in ('kungfu_b200', 'benchmarks', 'examples', 'tests', 'bench.py', '__graft_entry__.py'):
        p = os.path.join(ROOT, base)
        files = [p] if p.endswith('.py') else [os.path.join(r, f) for r, _, fs in os.walk(p) for f in fs if f.endswith('.py')]
        for path in files:
            for node in ast.walk(ast.parse(open(path).read())):
                if not isinstance(node, ast.Call) or any(isinstance(a, ast.Starred) for a in node.args) or node.keywords:
                    continue
                f = node.func
                if isinstance(f, ast.Attribute) and f.attr == 'call' and node.args and isinstance(node.args[0], ast.Constant) \
                        and isinstance(node.args[0].value, str) and node.args[0].value.startswith('kfb_'):
                    out.append((path, node.lineno, node.args[0].value, len(node.args) - 1))
                elif isinstance(f, ast.Attribute) and re.match(r'kfb?_[a-z0-9_]+$', f.attr):
                    out.append((path, node.lineno, f.attr, len(node.args)))
    return out


def test_every_call_site_passes_what_the_entry_point_takes():
    dev = _header_prototypes(os.path.join(ROOT, 'include', 'kungfu_b200.h'), 'kfb_')
    host = _header_prototypes(os.path.join(ROOT, 'include', 'kungfu_host.h'), 'kf_')
    sites = _call_sites()
    assert len(sites) > 150, len(sites)
    problems = []
    for path, line, name, nargs in sites:
        protos = dev if name.startswith('kfb_') else host
        if name not in protos:
            problems.append('%s:%d calls %s, which no header declares' % (os.path.relpath(path, ROOT), line, name))
        elif len(protos[name][1]) != nargs:
            problems.append('%s:%d passes %d arguments to %s, which takes %d' %
                            (os.path.relpath(path, ROOT), line, nargs, name, len(protos[name][1])))
    assert not problems, '\n'.join(problems)


def test_struct_layouts_and_constants_match_the_header(tmp_path):
    """kfb_comm / kfb_sgd_hyper as ctypes Structures against sizeof / offsetof printed by a C program that includes
    the header, and every constant the Python side restates (dtype / op / algo codes, flag bits, table sizes)."""
    import subprocess

    from kungfu_b200.ops import _cuda as K
    structs = {'kfb_comm': K.KfbComm, 'kfb_sgd_hyper': K.SgdHyper}
    rename = {'bias_corr1': 'bias_corr1', 'bias_corr2': 'bias_corr2'}
    consts = {'KFB_MAX_RANKS': K.MAX_RANKS, 'KFB_MAX_BLOCKS': K.MAX_BLOCKS, 'KFB_PAD_FLAG_WORDS': K.PAD_FLAG_WORDS,
              'KFB_PAD_MISC_WORD': K.PAD_MISC_WORD, 'KFB_PAD_BYTES': K.PAD_BYTES, 'KFB_STATE_EPOCH_WORDS': K.STATE_EPOCH_WORDS,
              'KFB_SCRATCH_FLOATS': K.SCRATCH_FLOATS, 'KFB_FUSE_SCRATCH_FLOATS': K.FUSE_SCRATCH_FLOATS,
              'KFB_VAR_WORK_WORDS': K.VAR_WORK_WORDS, 'KFB_F32': K.F32, 'KFB_BF16': K.BF16, 'KFB_F16': K.F16, 'KFB_I32': K.I32,
              'KFB_F64': K.F64, 'KFB_I64': K.I64, 'KFB_U8': K.U8, 'KFB_OP_SUM': K.OP_SUM, 'KFB_OP_MIN': K.OP_MIN,
              'KFB_OP_MAX': K.OP_MAX, 'KFB_OP_PROD': K.OP_PROD, 'KFB_ALGO_ONESHOT': K.ALGO_ONESHOT,
              'KFB_ALGO_TWOSHOT': K.ALGO_TWOSHOT, 'KFB_ALGO_NVLS': K.ALGO_NVLS, 'KFB_STEP_PRESIGNALED': K.STEP_PRESIGNALED,
              'KFB_STEP_NO_EXIT_BARRIER': K.STEP_NO_EXIT_BARRIER, 'KFB_STEP_PARTIAL_NORM': K.STEP_PARTIAL_NORM}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "kungfu_b200.h"', 'int main(void) {']
    for cname, st in structs.items():
        lines.append('  printf("sizeof %s %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in st._fields_:
            lines.append('  printf("offset %s.%s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, rename.get(fname, fname)))
    for c in consts:
        lines.append('  printf("const %s %%lld\\n", (long long)(%s));' % (c, c))
    lines += ['  return 0;', '}']
    src = tmp_path / 'probe.c'
    src.write_text('\n'.join(lines))
    exe = str(tmp_path / 'probe')
    subprocess.check_call(['gcc', '-std=c99', '-Wall', '-I', os.path.join(ROOT, 'include'), str(src), '-o', exe])
    got = {}
    for line in subprocess.run([exe], stdout=subprocess.PIPE, text=True, check=True).stdout.splitlines():
        kind, key, val = line.split()
        got[(kind, key)] = int(val)
    for cname, st in structs.items():
        assert got[('sizeof', cname)] == C.sizeof(st), (cname, got[('sizeof', cname)], C.sizeof(st))
        for fname, _ in st._fields_:
            assert got[('offset', '%s.%s' % (cname, fname))] == getattr(st, fname).offset, (cname, fname)
    for c, v in consts.items():
        assert got[('const', c)] == v, (c, got[('const', c)], v)
    # the host-side dtype codes (the reference's bit-packed encoding) against the runtime's enum
    from kungfu_b200 import _host as H
    text = open(os.path.join(ROOT, 'csrc', 'core', 'base.hpp')).read()
    codes = re.findall(r'\b([A-Z][A-Z0-9]+)\s*=\s*type_code\((\d+),\s*(\d+)\)', text)
    assert len(codes) >= 13, codes
    for name, cat, nbytes in codes:
        assert getattr(H, name) == (int(cat) << 16) | (int(nbytes) << 8) | 8, name
