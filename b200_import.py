"""Import helper: the package directory is named `gnss-sdr-new_b200` (not a Python identifier), so it is
loaded by path and registered as module `gnss_sdr_new_b200`."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.join(ROOT, "gnss-sdr-new_b200")


def load_package():
    name = "gnss_sdr_new_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG_DIR, "__init__.py"),
                                                  submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_oracle():
    """TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench cpu_baseline)."""
    odir = os.path.join(ROOT, "oracle")
    if odir not in sys.path:
        sys.path.insert(0, odir)
    import oracle  # noqa: E402
    return oracle
