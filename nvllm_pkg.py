"""Import helper: loads the package directory ``nano-vllm-go_b200/`` (hyphenated,
so not importable by name) as the module ``nano_vllm_go_b200``."""
import importlib
import importlib.util
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
_NAME = "nano_vllm_go_b200"


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    pkg_dir = os.path.join(_ROOT, "nano-vllm-go_b200")
    spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(pkg_dir, "__init__.py"), submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


def sub(name: str):
    """nvllm_pkg.sub('config') -> nano_vllm_go_b200.config"""
    load()
    return importlib.import_module(f"{_NAME}.{name}")
