"""Shim: TensorDict as a plain dict of tensors (ctx_manager.py:291-296)."""


class TensorDict(dict):
    def __init__(self, source=None, batch_size=None, **k):
        super().__init__(source or {})
        self.batch_size = batch_size
