"""Generic stub machinery: a module whose every attribute is a dummy class.

Dev/test tooling only. It lets the *reference* pybamm package (read-only at
/root/reference) be imported in a container that has none of its heavy
third-party dependencies (casadi, xarray, ...), so that model building,
parameterisation and discretisation can run and the discretised Symbol tree can
be lowered by pybamm_b200.frontend. Nothing here is used by the product path.
"""
import sys
import types


class _DummyMeta(type):
    def __getattr__(cls, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _make_dummy(f"{cls.__name__}.{name}")

    def __call__(cls, *args, **kwargs):
        return super().__call__(*args, **kwargs)


def _make_dummy(name):
    return _DummyMeta(name.split(".")[-1], (object,), {"__init__": lambda self, *a, **k: None})


class GenericStubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = _make_dummy(name)
        setattr(self, name, obj)
        return obj


def install(name):
    mod = GenericStubModule(name)
    mod.__path__ = []
    sys.modules[name] = mod
    return mod
