"""Shim: only OmegaConf.register_new_resolver is reached (ragen/utils.py:20-26)."""


class OmegaConf:
    @staticmethod
    def register_new_resolver(name, fn, **k):
        return None

    @staticmethod
    def create(obj):
        raise NotImplementedError("use oracle.cfg.AttrDict")
