"""gymtorch stand-in: tensors are already torch tensors, wrap/unwrap are identity."""


def wrap_tensor(t):
    return t


def unwrap_tensor(t):
    return t
