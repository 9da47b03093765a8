"""garfield_b200.ops — Python drivers of the hand-written sm_100a kernels.

gar      aggregation rules (pairwise_select + coordwise kernels, fused SGD epilogue)
optim    flat-arena optimizers (sgd/adam/adamw/rmsprop/adagrad)
attacks  Byzantine attack generation
"""

from . import _lib  # noqa: F401
