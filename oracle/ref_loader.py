"""TEST INFRASTRUCTURE ONLY -- loader that executes the UNMODIFIED reference source.

Nothing in the product package may import this file.  Allowed importers: tests/, oracle/gen_golden*.py,
__graft_entry__.smoke() and bench.py's `cpu_baseline` / `--impl reference` legs.

Where the reference comes from: /root/reference in the build container; on the GPU box (where that path does not
exist) the byte-for-byte staged copy oracle/_ref/ (git-ignored, shipped with the snapshot; oracle/stage_ref.py
writes it during build()).  `source()` says which one is in use.

What it does
------------
`import dynamo` fails in this image (anndata, matplotlib, pynndescent, numdifftools, ... are not
installed and there is no network).  The numeric hot-path files themselves only need numpy / scipy /
sklearn / pandas / tqdm, so this loader

  1. registers a `sys.meta_path` finder that serves *permissive stub modules* for the missing
     third-party roots (they are touched only for logging / plotting configuration, never for
     arithmetic), and
  2. registers *bare* package objects for `dynamo`, `dynamo.tools`, `dynamo.vectorfield`, ... whose
     `__path__` points into /root/reference/dynamo/*, so the heavy `__init__.py` files are skipped
     and `importlib.import_module("dynamo.vectorfield.scVectorField")` executes the reference file
     itself.

Reference files that end up executing (all unmodified, read from /root/reference):
  dynamo/vectorfield/scVectorField.py, utils.py, vector_calculus.py, VectorField.py, rank_vf.py,
  dynamo/tools/utils.py, sampling.py, connectivity.py, dynamo/simulation/ODE.py, dynamo/dynamo_logger.py
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.machinery
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
STAGED_ROOT = os.path.join(_HERE, "_ref")


def _pick_root():
    env = os.environ.get("DVF_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/dynamo/vectorfield"):
        return "/root/reference"
    return STAGED_ROOT


REFERENCE_ROOT = _pick_root()

# third-party roots that are absent here and only used for non-numeric side effects
_STUB_ROOTS = {
    "anndata", "matplotlib", "mpl_toolkits", "pynndescent", "numdifftools", "loompy", "umap",
    "seaborn", "colorcet", "cycler", "igraph", "leidenalg", "statsmodels", "patsy", "mudata",
    "session_info", "adjustText", "datacache", "get_version", "openpyxl", "hnswlib", "louvain",
    "cdlib", "trimap", "fitsne", "pacmap", "typing_extensions_stub", "scanpy", "KDEpy", "numba_stub",
    "dynode", "mygene", "gseapy", "pysal", "nxviz", "hiveplotlib", "yt", "vaex", "networkit",
}


class _Anything:
    """Catch-all object: callable, subscriptable, iterable (empty), comparable, usable as a base class."""

    def __init__(self, name="stub"):
        object.__setattr__(self, "_name", name)

    def __call__(self, *a, **k):
        # used as decorator -> return the function unchanged
        if len(a) == 1 and not k and callable(a[0]) and not isinstance(a[0], _Anything):
            return a[0]
        return _Anything(self._name + "()")

    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Anything(self._name + "." + item)

    def __setattr__(self, k, v):
        pass

    def __getitem__(self, k):
        return _Anything(self._name + "[]")

    def __setitem__(self, k, v):
        pass

    def __delitem__(self, k):
        pass

    def __contains__(self, k):
        return False

    def __iter__(self):
        return iter(())

    def __len__(self):
        return 0

    def __bool__(self):
        return False

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def __mro_entries__(self, bases):
        return (object,)

    def __repr__(self):
        return "<stub %s>" % self._name

    def _binop(self, *a):
        return _Anything(self._name)

    __add__ = __radd__ = __sub__ = __rsub__ = __mul__ = __rmul__ = __truediv__ = __rtruediv__ = _binop
    __floordiv__ = __mod__ = __pow__ = __and__ = __or__ = __neg__ = __pos__ = _binop

    def __lt__(self, o):
        return False

    __le__ = __gt__ = __ge__ = __lt__

    def __eq__(self, o):
        return self is o

    def __hash__(self):
        return id(self)


class _StubModule(types.ModuleType):
    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Anything(self.__name__ + "." + item)


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        root = fullname.split(".")[0]
        if root in _STUB_ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        if spec.name == "anndata":
            # a real (empty) class so isinstance(..., AnnData) and type hints work
            m.AnnData = type("AnnData", (), {})
        return m

    def exec_module(self, module):
        pass


_BARE_PACKAGES = [
    "dynamo", "dynamo.tools", "dynamo.vectorfield", "dynamo.simulation", "dynamo.preprocessing",
    "dynamo.prediction", "dynamo.plot", "dynamo.estimation", "dynamo.external", "dynamo.movie",
]

_installed = False


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "dynamo", "vectorfield"))


def source() -> str:
    """'reference' (the tree itself), '_ref' (the staged byte-for-byte copy) or 'absent'."""
    if not available():
        return "absent"
    return "_ref" if os.path.abspath(REFERENCE_ROOT) == os.path.abspath(STAGED_ROOT) else "reference"


def verify_staged() -> int:
    """Check every staged file against the sha256 recorded when it was copied; returns the number of files."""
    import hashlib
    import json

    with open(os.path.join(STAGED_ROOT, "MANIFEST.json")) as f:
        man = json.load(f)["files"]
    for rel_path, digest in man.items():
        with open(os.path.join(STAGED_ROOT, rel_path), "rb") as f:
            if hashlib.sha256(f.read()).hexdigest() != digest:
                raise RuntimeError("staged reference file %s was modified" % rel_path)
    return len(man)


def install():
    """Idempotent.  After this, `importlib.import_module('dynamo.vectorfield.utils')` runs the reference file."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
    # only stub what is really missing
    for root in list(_STUB_ROOTS):
        try:
            if importlib.util.find_spec(root) is not None:
                _STUB_ROOTS.discard(root)
        except (ImportError, ValueError):
            pass
    sys.meta_path.append(_StubFinder())
    import numpy.matlib  # noqa: F401  (reference utils.py:204 uses np.matlib without importing it)

    for name in _BARE_PACKAGES:
        if name in sys.modules:
            continue
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, *name.split("."))]
        pkg.__package__ = name
        sys.modules[name] = pkg
        if "." in name:
            parent, _, child = name.rpartition(".")
            setattr(sys.modules[parent], child, pkg)
    _installed = True


def load(modname: str):
    """Import a reference module by dotted name, e.g. 'dynamo.vectorfield.scVectorField'."""
    install()
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # reference files carry invalid-escape SyntaxWarnings
        _quiet_logger()
        mod = importlib.import_module(modname)
    return mod


_logger_quiet = False


def _quiet_logger():
    """Silence the reference's stdout logger (logging only; arithmetic untouched).

    reference dynamo_logger.py:62-99 sets every namespaced logger to INFO on construction, so the
    constructor is wrapped to lower it again."""
    global _logger_quiet
    if _logger_quiet or os.environ.get("DVF_REF_VERBOSE"):
        return
    import logging

    lm = importlib.import_module("dynamo.dynamo_logger")
    orig_init = lm.Logger.__init__

    def quiet_init(self, *a, **k):
        orig_init(self, *a, **k)
        self.logger.setLevel(logging.CRITICAL + 100)

    lm.Logger.__init__ = quiet_init
    for lg in (lm.LoggerManager.main_logger, lm.LoggerManager.temp_timer_logger):
        lg.logger.setLevel(logging.CRITICAL + 100)
    _logger_quiet = True


def hot_path():
    """Return (scVectorField, vf_utils, tools_utils) reference modules."""
    vfu = load("dynamo.vectorfield.utils")
    sc = load("dynamo.vectorfield.scVectorField")
    tu = load("dynamo.tools.utils")
    return sc, vfu, tu
