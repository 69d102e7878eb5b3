"""Package shell so that `from verl import DataProto` and `verl.utils.torch_functional`
resolve; the real algorithm files are loaded by path from /root/reference/verl/verl by
oracle/ref_loader.py."""


class DataProto:
    def __init__(self, batch=None, non_tensor_batch=None, meta_info=None):
        self.batch = batch
        self.non_tensor_batch = non_tensor_batch if non_tensor_batch is not None else {}
        self.meta_info = meta_info if meta_info is not None else {}
