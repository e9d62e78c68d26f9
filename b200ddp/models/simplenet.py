"""SimpleNet — the reference's model (reference ``train.py:32-50``).

``Flatten -> Linear(784,256) -> ReLU -> Dropout(0.2) -> Linear(256,256) -> ReLU
-> Dropout(0.2) -> Linear(256,10)``; 269,322 parameters.  The ``nn.Sequential``
positions (0/3/6) are kept so ``state_dict`` keys are ``layers.0.weight`` …
``layers.6.bias`` and checkpoints cross-load with the reference.

The module itself is plain ``torch.nn`` so it runs anywhere (CPU/gloo plumbing
config).  On a B200 the engine does not call ``forward`` for training: it reads
``mlp_spec()`` and runs the whole step (three tcgen05 GEMMs forward with fused
bias/ReLU/dropout epilogues, fused cross-entropy, five GEMMs backward writing
straight into the gradient arena, fused all-reduce + Adam) as one captured
CUDA graph; see ``b200ddp/parallel/mlp_program.py``.
"""
from __future__ import annotations

from typing import List, Tuple

import torch.nn as nn


class SimpleNet(nn.Module):
    def __init__(self, input_size: int = 784, hidden_size: int = 256, num_classes: int = 10,
                 dropout: float = 0.2):
        super().__init__()
        self.input_size = input_size
        self.hidden_size = hidden_size
        self.num_classes = num_classes
        self.dropout_p = dropout
        self.flatten = nn.Flatten()
        self.layers = nn.Sequential(
            nn.Linear(input_size, hidden_size),
            nn.ReLU(),
            nn.Dropout(dropout),
            nn.Linear(hidden_size, hidden_size),
            nn.ReLU(),
            nn.Dropout(dropout),
            nn.Linear(hidden_size, num_classes),
        )

    def forward(self, x):
        x = self.flatten(x)
        return self.layers(x)

    def mlp_spec(self) -> List[Tuple[nn.Linear, bool, float]]:
        """(linear, relu_after, dropout_p_after) per layer, for the fused step program."""
        return [
            (self.layers[0], True, self.dropout_p),
            (self.layers[3], True, self.dropout_p),
            (self.layers[6], False, 0.0),
        ]
