"""Import shim that loads the UNMODIFIED reference package: `/root/reference/src` in the build container,
else `baseline/_ref` (the copy `baseline/install_ref.py` makes; it travels to the GPU box).

TEST / BASELINE TOOLING ONLY.  It is used by `tests/golden/make_golden.py` (golden-vector generation), by the
drop-in tests (which need a real `bloqade.analog` to drive `program.bloqade.python().run(...)` on the B200
routine and to compare it with the reference's own routine in the same process), and by `bench.py`'s
reference arm and `.run()` end-to-end leg.  Nothing in the product package imports this file.

The reference's hot path (emulate/*) never touches plotting / cloud SDKs, but `import bloqade.analog`
pulls them in transitively.  They are absent from this image and there is no network, so they are
replaced by inert stand-ins:

* plotext, bokeh.*, braket.*, requests_sigv4 -> stub modules whose attributes are dummy classes
* simplejson -> stdlib json (only `dumps(use_decimal=...)` differs; not on the hot path)
* plum -> a first-match overload dispatcher driven by beartype (needed by
  `StateVector.local_trace`, state_vector.py:88-202, and builder/field.py:43-62)
* importlib.metadata.version("bloqade-analog") -> the pyproject version
"""
import functools
import importlib.machinery
import importlib.metadata as _md
import inspect
import json as _json
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_SRC = os.environ.get("BQA_REFERENCE_SRC", "/root/reference/src")  # override: test the copy here
if not os.path.isdir(os.path.join(REFERENCE_SRC, "bloqade", "analog")):
    REFERENCE_SRC = os.path.join(_HERE, "_ref")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "bloqade", "analog"))


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything


def _stub(name):
    mod = _StubModule(name)
    mod.__path__ = []
    mod.__spec__ = importlib.machinery.ModuleSpec(name, None)
    sys.modules[name] = mod
    if "." in name:
        parent, child = name.rsplit(".", 1)
        if parent in sys.modules:
            setattr(sys.modules[parent], child, mod)
    return mod


_STUBBED = [
    "plotext",
    "bokeh", "bokeh.io", "bokeh.models", "bokeh.layouts", "bokeh.palettes", "bokeh.plotting",
    "bokeh.models.widgets",
    "braket", "braket.ir", "braket.ir.ahs", "braket.timings", "braket.ahs", "braket.ahs.field",
    "braket.ahs.pattern", "braket.task_result", "braket.ahs.driving_field",
    "braket.ahs.shifting_field", "braket.ahs.atom_arrangement",
    "braket.ahs.analog_hamiltonian_simulation", "braket.aws", "braket.devices",
    "requests_sigv4",
]

_installed = False


def _install_plum():
    from beartype.door import is_bearable

    plum = _stub("plum")
    registry = {}

    def register(fn):
        key = (fn.__module__, fn.__qualname__)
        impls = registry.setdefault(key, [])
        body_is_placeholder = bool(fn.__doc__) and len(fn.__code__.co_code) <= 8
        if not body_is_placeholder:
            impls.append((inspect.signature(fn), fn.__annotations__, fn))

        @functools.wraps(fn)
        def call(*args, **kwargs):
            for sig, ann, impl in registry[key]:
                try:
                    bound = sig.bind(*args, **kwargs)
                except TypeError:
                    continue
                bound.apply_defaults()
                ok = True
                for pname, val in bound.arguments.items():
                    tp = ann.get(pname)
                    if tp is None or isinstance(tp, str):
                        continue
                    try:
                        if not is_bearable(val, tp):
                            ok = False
                            break
                    except Exception:
                        pass
                if ok:
                    return impl(*args, **kwargs)
            raise TypeError(f"no overload of {key} matches {args[1:]} {kwargs}")

        return call

    plum.overload = register
    plum.dispatch = register


def install():
    """Make `import bloqade.analog` resolve to the reference tree.  Idempotent."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_SRC}")
    for name in _STUBBED:
        if name not in sys.modules:
            _stub(name)
    from pydantic.v1 import BaseModel

    class _AhsProgram(BaseModel):
        class Config:
            extra = "allow"

    sys.modules["braket.ir.ahs"].Program = _AhsProgram
    sys.modules.setdefault("simplejson", _json)
    if "plum" not in sys.modules:
        _install_plum()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    orig_version = _md.version

    def version(name):
        if name == "bloqade-analog":
            return "0.17.0-dev"
        return orig_version(name)

    _md.version = version
    import bloqade.analog  # noqa: F401

    _installed = True
