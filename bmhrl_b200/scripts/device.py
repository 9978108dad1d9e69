"""scripts/device.py:3-10 of the reference: cfg.device -> torch.device (selecting device_ids[0] for CUDA)."""
import torch


def get_device(cfg):
    if torch.cuda.is_available() and str(getattr(cfg, "device", "cuda")).startswith("cuda"):
        ids = getattr(cfg, "device_ids", None) or [torch.cuda.current_device()]
        torch.cuda.set_device(ids[0])
        return torch.device(f"cuda:{ids[0]}")
    return torch.device("cpu")
