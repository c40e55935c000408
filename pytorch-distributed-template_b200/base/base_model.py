"""BaseModel: the plug-in base class every `arch.type` derives from
(reference: base/base_model.py:6-25)."""
from abc import abstractmethod

import torch.nn as nn


class BaseModel(nn.Module):
    @abstractmethod
    def forward(self, *inputs):
        raise NotImplementedError

    def num_trainable(self):
        return sum(p.numel() for p in self.parameters() if p.requires_grad)

    def __str__(self):
        # same trailer the reference prints under the module tree
        return "{}\nTrainable parameters: {}".format(super().__str__(), self.num_trainable())
