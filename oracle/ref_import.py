"""Import the UNMODIFIED reference miniapps from /root/reference (build container only).

TEST INFRASTRUCTURE (see oracle/__init__.py).  The reference modules import
``matplotlib`` (fluence_mini_app.py:9-10) and call ``coloredlogs.install``
(utils/logging_utils.py:65); neither package is installed here, so two inert
stub modules are injected into ``sys.modules`` before loading.  Nothing else of
the reference is touched: ``run``, ``run_pc`` and ``sourceTerm`` execute the
reference's own code.

``/root/reference`` does not exist on the GPU box, so this module is used only
(a) by ``oracle/make_golden.py`` to generate the committed fixtures and (b) by
CPU tests that skip when the reference is absent.
"""
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("NRW_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "fluence_mini_app.py"))


def _install_stubs():
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except ImportError:
            m = types.ModuleType("matplotlib")
            m.use = lambda *a, **k: None
            sys.modules["matplotlib"] = m
    if "coloredlogs" not in sys.modules:
        try:
            import coloredlogs  # noqa: F401
        except ImportError:
            c = types.ModuleType("coloredlogs")
            c.install = lambda *a, **k: None
            sys.modules["coloredlogs"] = c


def _load(name: str, alias: str):
    """Load /root/reference/<name>.py under the module name ``alias``.

    The reference does ``import utils.logging_utils as utils``; that must resolve
    to the reference's own ``utils`` package, so the reference root is put at
    the front of ``sys.path`` for the duration of the load and the reference's
    ``utils`` is evicted afterwards so that it cannot shadow the product's.
    """
    if not reference_available():
        raise FileNotFoundError(
            "reference not present at %s (expected on the GPU box)" % REFERENCE_ROOT)
    _install_stubs()
    saved_path = list(sys.path)
    saved_utils = {k: v for k, v in sys.modules.items()
                   if k == "utils" or k.startswith("utils.")}
    for k in saved_utils:
        del sys.modules[k]
    import signal
    saved_handlers = (signal.getsignal(signal.SIGTERM), signal.getsignal(signal.SIGINT))
    try:
        sys.path.insert(0, REFERENCE_ROOT)
        spec = importlib.util.spec_from_file_location(
            alias, os.path.join(REFERENCE_ROOT, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path[:] = saved_path
        for k in [k for k in sys.modules if k == "utils" or k.startswith("utils.")]:
            del sys.modules[k]
        sys.modules.update(saved_utils)
        # the reference installs process-wide signal handlers at import
        # (fluence_mini_app.py:25-26); undo that for the test process.
        try:
            signal.signal(signal.SIGTERM, saved_handlers[0])
            signal.signal(signal.SIGINT, saved_handlers[1])
        except (ValueError, TypeError):
            pass
    return mod


_cache = {}


def fluence():
    """The reference's fluence_mini_app module (run, run_pc, sourceTerm ...)."""
    if "fluence" not in _cache:
        _cache["fluence"] = _load("fluence_mini_app", "_nrw_ref_fluence")
    return _cache["fluence"]


def simple_flux():
    """The reference's simple_flux_miniapp module."""
    if "simple" not in _cache:
        _cache["simple"] = _load("simple_flux_miniapp", "_nrw_ref_simple_flux")
    return _cache["simple"]


def fluence_default_args(**over):
    """argparse defaults of fluence_mini_app.py:493-512 as the dict ``run`` takes."""
    import logging
    a = dict(run_mode="pc", num_walks=100, timestep=0.01, anglestep=1.0 / 15.0,
             velocity=200, SigS=0.15, SigA=0, interval_length=1,
             neural_timesteps=30000, loglevel=logging.WARNING,
             matrix_filename="curmat.csv", results_filename="rwSol.csv",
             fugu_backend="snn", plot_results=False,
             log_file="~/.miniapps/fluence_miniapp.log", runclean=0,
             rand_seed=-1, use_sinks=False)
    a.update(over)
    return a


def simple_flux_default_args(**over):
    """argparse defaults of simple_flux_miniapp.py:392-409."""
    import logging
    a = dict(run_mode="pc", num_walks=500, timestep=0.005, init_pos=5.0,
             init_neg=3.0, SigS=8.0, SigA=2.0, max_time=5.0,
             neural_timesteps=500000, loglevel=logging.WARNING,
             matrix_filename="curmat.csv", results_filename="rwSol.csv",
             fugu_backend="snn", plot_results=False,
             log_file="~/.miniapps/simple_flux_miniapp.log",
             rand_seed=-1, use_sinks=False)
    a.update(over)
    return a


# ------------------------------------------------------------------------------------------
# the "next" surface: loihirw.rw_analysis / rw_core, densityrw.DensityRW
# ------------------------------------------------------------------------------------------

_STUB_MODULES = ("nxsdk", "nxsdk.api", "nxsdk.api.n2a", "nxsdk.api.enums", "nxsdk.api.enums.api_enums",
                 "nxsdk.graph", "nxsdk.graph.monitor", "nxsdk.graph.monitor.probes",
                 "fugu", "fugu.bricks", "fugu.scaffold", "fugu.backends")


def _load_packaged(relpath: str, alias: str, evict=("densityrw", "loihirw")):
    """Load one file of a reference sub-package with inert stand-ins for the proprietary / absent
    imports it makes at module level (NxSDK, Fugu).  Only pure-NumPy functions of those files are
    ever called (rw_analysis.process_outputs*, rw_core.get_probCircuitValues,
    DensityRW.load_transitions / build_test_graph)."""
    if not reference_available():
        raise FileNotFoundError("reference not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    added = []
    for name in _STUB_MODULES:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
            added.append(name)

    class _Inert:
        def __init__(self, *a, **k):
            pass
    sys.modules["fugu"].input_coding_types = []
    for mod, attrs in (("fugu.bricks", ("Brick", "Vector_Input")), ("fugu.scaffold", ("Scaffold",)),
                       ("fugu.backends", ("snn_Backend",))):
        for a in attrs:
            setattr(sys.modules[mod], a, _Inert)
    saved_path = list(sys.path)
    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] in evict}
    for k in saved:
        del sys.modules[k]
    import numpy as _np
    had_nan = hasattr(_np, "NaN")
    if not had_nan:
        _np.NaN = _np.nan                        # SpikingPDE.py:367,385 predates numpy 2
    try:
        sys.path.insert(0, REFERENCE_ROOT)
        spec = importlib.util.spec_from_file_location(alias, os.path.join(REFERENCE_ROOT, relpath))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path[:] = saved_path
        for k in [k for k in sys.modules if k.split(".")[0] in evict]:
            del sys.modules[k]
        sys.modules.update(saved)
        for name in added:
            sys.modules.pop(name, None)
        if not had_nan:
            try:
                del _np.NaN
            except AttributeError:
                pass
    return mod


def rw_analysis():
    """The reference's loihirw/rw_analysis.py (pure NumPy apart from its NxSDK star-import)."""
    if "rw_analysis" not in _cache:
        _cache["rw_analysis"] = _load_packaged("loihirw/rw_analysis.py", "_nrw_ref_rw_analysis")
    return _cache["rw_analysis"]


def rw_core():
    """The reference's loihirw/rw_core.py (only get_probCircuitValues is NxSDK-free)."""
    if "rw_core" not in _cache:
        _cache["rw_core"] = _load_packaged("loihirw/rw_core.py", "_nrw_ref_rw_core")
    return _cache["rw_core"]


def density_rw():
    """The reference's densityrw/DensityRW.py (load_transitions, build_test_graph)."""
    if "density_rw" not in _cache:
        _cache["density_rw"] = _load_packaged("densityrw/DensityRW.py", "_nrw_ref_density_rw")
    return _cache["density_rw"]
