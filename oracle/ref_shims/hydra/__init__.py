"""Shim: hydra.main as an identity decorator (es_manager.py:263, ctx_manager.py:362)."""


def main(*a, **k):
    def deco(fn):
        return fn
    return deco
