"""In-process loader for the UNMODIFIED reference (djones1040/PythonPhot).

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file.
It exists so that (a) the C restatement in ``oracle/bphot_oracle.c`` can be
pinned against the reference itself, and (b) ``tests/golden/gen_golden.py``
can generate the committed golden vectors.  It only works in the authoring
container, where the reference checkout is mounted at ``/root/reference``;
the GPU box has no such directory and nothing there calls this module.

The reference package cannot be imported normally here (its ``__init__``
pulls in astropy, PythonPhot/__init__.py:10) and three of its modules use
numpy aliases removed in numpy 2 (aper.py:24 ``np.asfarray``; getpsf.py:412
``np.int``).  We therefore register a synthetic package object whose
``__path__`` is the reference directory and patch the missing numpy names.
No reference file is edited or copied.

Two semantic modes:

* ``py3`` -- the modules exactly as Python 3 runs them.
* ``py2`` -- every ``a / b`` is rewritten (on the AST, in memory) to Python 2
  division: floor division when both operands are integers.  The reference
  was written for Python 2.7; the only hot-path expression that differs is
  ``yy = nbox/nx[i]`` in aper.py:248.
"""
import ast
import contextlib
import importlib
import io
import os
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("BPHOT_REFERENCE_ROOT", "/root/reference")
REF_PKG_DIR = os.path.join(REF_ROOT, "PythonPhot")


def available():
    return os.path.isdir(REF_PKG_DIR)


def _patch_numpy():
    if not hasattr(np, "asfarray"):
        np.asfarray = lambda a, dtype=float: np.asarray(a, dtype=dtype)
    for name, typ in (("int", int), ("float", float)):
        if name not in np.__dict__:
            setattr(np, name, typ)


class _HDU:
    def __init__(self):
        self.header = {}
        self.data = None

    def writeto(self, *a, **k):
        pass


def _stub_pyfits():
    for name in ("pyfits",):
        if name not in sys.modules:
            stub = types.ModuleType(name)
            stub.PrimaryHDU = _HDU
            stub.getdata = lambda *a, **k: None
            stub.getheader = lambda *a, **k: {}
            sys.modules[name] = stub


def _py2div(a, b):
    def _is_int(v):
        if isinstance(v, (bool, np.bool_)):
            return False
        if isinstance(v, (int, np.integer)):
            return True
        if isinstance(v, range):
            return True
        if isinstance(v, np.ndarray) and np.issubdtype(v.dtype, np.integer):
            return True
        return False

    if _is_int(a) and _is_int(b):
        if isinstance(a, range):
            a = np.asarray(a)
        if isinstance(b, range):
            b = np.asarray(b)
        return a // b
    if isinstance(a, range):
        a = np.asarray(a)
    if isinstance(b, range):
        b = np.asarray(b)
    return a / b


class _DivRewriter(ast.NodeTransformer):
    def visit_BinOp(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):
            return ast.copy_location(
                ast.Call(func=ast.Name(id="_py2div", ctx=ast.Load()),
                         args=[node.left, node.right], keywords=[]), node)
        return node


_PKG = {"py3": "PythonPhot", "py2": "PythonPhot_py2"}


def _ensure_pkg(mode):
    name = _PKG[mode]
    if name not in sys.modules:
        pkg = types.ModuleType(name)
        pkg.__path__ = [REF_PKG_DIR]
        sys.modules[name] = pkg
    return name


def load(modname, mode="py3"):
    """Return reference module ``modname`` (e.g. 'aper') under ``mode`` semantics."""
    if not available():
        raise RuntimeError("reference checkout not found at %s" % REF_PKG_DIR)
    _patch_numpy()
    _stub_pyfits()
    pkgname = _ensure_pkg(mode)
    full = pkgname + "." + modname
    if full in sys.modules:
        return sys.modules[full]
    if mode == "py3":
        with contextlib.redirect_stdout(io.StringIO()):
            return importlib.import_module(full)
    # py2: exec the AST-rewritten source into a fresh module of the py2 package
    path = os.path.join(REF_PKG_DIR, modname + ".py")
    with open(path) as fh:
        tree = ast.parse(fh.read(), filename=path)
    # resolve "from . import a, b" first so relative imports find py2 siblings
    for node in ast.walk(tree):
        if isinstance(node, ast.ImportFrom) and node.level == 1 and node.module is None:
            for alias in node.names:
                load(alias.name, mode)
    tree = ast.fix_missing_locations(_DivRewriter().visit(tree))
    mod = types.ModuleType(full)
    mod.__package__ = pkgname
    mod.__file__ = path
    mod.__dict__["_py2div"] = _py2div
    sys.modules[full] = mod
    setattr(sys.modules[pkgname], modname, mod)
    exec(compile(tree, path, "exec"), mod.__dict__)
    return mod


@contextlib.contextmanager
def quiet():
    """The reference prints unconditionally (mmm.py:111-256, dao_value.py:73)."""
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        yield buf
