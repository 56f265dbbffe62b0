import torch


def wrap_phase_err(a, b):
    d = (a - b).double()
    return torch.atan2(torch.sin(d), torch.cos(d)).abs()
