"""Make the reference pybamm importable in this container (dev/test tooling).

Usage:  import tools.pybamm_env as env; pybamm = env.load()

The reference tree is read-only and is NOT available on the GPU box; only the
frozen model bundles produced with it (see tools/freeze_models.py) travel.
Recipe follows SURVEY.md §8c.
"""
import os
import sys
import types

REFERENCE = os.environ.get("PYBAMM_REFERENCE", "/root/reference")
_SRC = os.path.join(REFERENCE, "packages", "pybamm", "src")


def available():
    return os.path.isdir(os.path.join(_SRC, "pybamm"))


def _install_stubs():
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pybamm_stubs")
    if here not in sys.path:
        sys.path.insert(0, here)
    import _generic

    for name in ("xarray", "pooch", "black"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                _generic.install(name)
    # casadi: generic stub; binary_operators does isinstance(x, casadi.casadi.MX)
    try:
        import casadi  # noqa: F401
    except Exception:
        cas = _generic.install("casadi")
        cas.casadi = cas
        sys.modules["casadi.casadi"] = cas
    # posthog
    try:
        import posthog  # noqa: F401
    except Exception:
        ph = types.ModuleType("posthog")

        class Posthog:
            def __init__(self, *a, **k):
                self.disabled = True

            def capture(self, *a, **k):
                pass

        ph.Posthog = Posthog
        sys.modules["posthog"] = ph
    # pybammsolvers + idaklu
    try:
        import pybammsolvers  # noqa: F401
    except Exception:
        ps = _generic.install("pybammsolvers")
        idk = _generic.install("pybammsolvers.idaklu")
        ps.idaklu = idk
    # anytree: only the two iterators are used (Symbol.pre_order / post_order)
    try:
        import anytree  # noqa: F401
    except Exception:
        at = types.ModuleType("anytree")

        class PreOrderIter:
            def __init__(self, node, filter_=None, stop=None, maxlevel=None):
                self.node = node
                self.filter_ = filter_

            def __iter__(self):
                stack = [self.node]
                while stack:
                    n = stack.pop()
                    if self.filter_ is None or self.filter_(n):
                        yield n
                    stack.extend(reversed(list(n.children)))

        class PostOrderIter:
            def __init__(self, node, filter_=None, stop=None, maxlevel=None):
                self.node = node
                self.filter_ = filter_

            def __iter__(self):
                def rec(n):
                    for c in n.children:
                        yield from rec(c)
                    if self.filter_ is None or self.filter_(n):
                        yield n

                return rec(self.node)

        class NodeMixin:
            pass

        at.PreOrderIter = PreOrderIter
        at.PostOrderIter = PostOrderIter
        at.NodeMixin = NodeMixin
        exporter = types.ModuleType("anytree.exporter")

        class DotExporter:
            def __init__(self, *a, **k):
                pass

        exporter.DotExporter = DotExporter
        at.exporter = exporter
        sys.modules["anytree"] = at
        sys.modules["anytree.exporter"] = exporter
    # hatch-vcs version file is absent in a source checkout
    if "pybamm._version" not in sys.modules:
        v = types.ModuleType("pybamm._version")
        v.__version__ = "0.0.0+reference"
        v.version = v.__version__
        sys.modules["pybamm._version"] = v


_pybamm = None


def load():
    """Import and return the reference pybamm module (telemetry disabled)."""
    global _pybamm
    if _pybamm is not None:
        return _pybamm
    if not available():
        raise ImportError(f"reference pybamm not found under {REFERENCE}")
    os.environ.setdefault("PYBAMM_DISABLE_TELEMETRY", "true")
    _install_stubs()
    if _SRC not in sys.path:
        sys.path.insert(0, _SRC)
    import pybamm

    _pybamm = pybamm
    return pybamm
