"""Forward pass of every model family on random input (the reference's ``__main__`` blocks in models/*.py):
    python -m dba_mod_b200.models
"""
import torch

from . import BatchedNet, initial_state

for kind, shape in [("cifar", (3, 32, 32)), ("tiny-imagenet-200", (3, 64, 64)), ("mnist", (1, 28, 28)), ("loan", (91,))]:
    spec, flat, cnt = initial_state(kind)
    net = BatchedNet(kind, spec)
    x = torch.randn(1, *shape)
    y = net.forward(flat[None, :spec.Dp], flat[None, spec.Dp:], None, x, training=False, shared_input=True)
    print(kind, "params", spec.Dp, "state_dict elems", spec.D, "output", tuple(y.shape[1:]))
