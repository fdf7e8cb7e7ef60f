"""`deepspeed.accelerator.get_accelerator()` as train.py:12,61 uses it: set_device."""
import torch


class _CudaAccelerator:
    def device_name(self, index=None):
        return "cuda" if index is None else f"cuda:{index}"

    def set_device(self, index):
        torch.cuda.set_device(index)

    def current_device(self):
        return torch.cuda.current_device()

    def device_count(self):
        return torch.cuda.device_count()

    def is_available(self):
        return torch.cuda.is_available()

    def synchronize(self, index=None):
        torch.cuda.synchronize(index)

    def communication_backend_name(self):
        return "nccl"


_ACC = _CudaAccelerator()


def get_accelerator():
    return _ACC
