"""TEST INFRASTRUCTURE — loads the UNMODIFIED reference from /root/reference in this container.

Only `tests/` and `tests/golden/make_golden.py` use this.  It never runs on the GPU box
(`/root/reference` does not exist there): callers must check `available()` first.

What it does
  * puts `oracle/ref_shims` (stand-ins for absent third-party packages, see its README) and
    `/root/reference` on `sys.path`, so `ragen.env.*`, `ragen.llm_agent.es_manager`,
    `ragen.llm_agent.ctx_manager` import unmodified;
  * loads `verl/verl/utils/torch_functional.py`, `verl/verl/trainer/ppo/core_algos.py` by file path
    under their canonical module names, so `ragen/trainer/core_algos.py` imports unmodified.
"""
import importlib
import importlib.util
import os
import sys

REF_ROOT = os.environ.get("RAGEN_REFERENCE_ROOT", "/root/reference")
_SHIMS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_shims")
_loaded = {}


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "ragen"))


def _ensure_paths():
    if not available():
        raise RuntimeError(f"reference not present at {REF_ROOT}")
    for p in (_SHIMS, REF_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    # shims must win over anything else of the same name
    if sys.path.index(_SHIMS) > sys.path.index(REF_ROOT):
        sys.path.remove(_SHIMS)
        sys.path.insert(0, _SHIMS)


def _load_by_path(mod_name: str, rel_path: str):
    if mod_name in sys.modules and getattr(sys.modules[mod_name], "__ref_loaded__", False):
        return sys.modules[mod_name]
    path = os.path.join(REF_ROOT, rel_path)
    spec = importlib.util.spec_from_file_location(mod_name, path)
    mod = importlib.util.module_from_spec(spec)
    mod.__ref_loaded__ = True
    sys.modules[mod_name] = mod
    spec.loader.exec_module(mod)
    parent, _, leaf = mod_name.rpartition(".")
    if parent and parent in sys.modules:
        setattr(sys.modules[parent], leaf, mod)
    return mod


def module(name: str):
    """Import a `ragen.*` module of the reference unmodified."""
    _ensure_paths()
    if name not in _loaded:
        _loaded[name] = importlib.import_module(name)
    return _loaded[name]


def sokoban_utils():
    return module("ragen.env.sokoban.utils")


def sokoban_env():
    return module("ragen.env.sokoban.env")


def frozen_lake_env():
    return module("ragen.env.frozen_lake.env")


def bandit_env():
    return module("ragen.env.bandit.env")


def countdown_env():
    return module("ragen.env.countdown.env")


def es_manager():
    return module("ragen.llm_agent.es_manager")


def ctx_manager():
    return module("ragen.llm_agent.ctx_manager")


def torch_functional():
    _ensure_paths()
    import verl  # noqa: F401  (shim shell)
    import verl.utils  # noqa: F401
    return _load_by_path("verl.utils.torch_functional", "verl/verl/utils/torch_functional.py")


def verl_core_algos():
    torch_functional()
    import verl.trainer.ppo  # noqa: F401
    return _load_by_path("verl.trainer.ppo.core_algos", "verl/verl/trainer/ppo/core_algos.py")


def ragen_core_algos():
    verl_core_algos()
    return _load_by_path("ragen.trainer.core_algos", "ragen/trainer/core_algos.py")
