import collections.abc
from itertools import repeat
import torch.nn as nn
from torch.nn.init import trunc_normal_ as _tn


def to_2tuple(x):
    if isinstance(x, collections.abc.Iterable) and not isinstance(x, str):
        return tuple(x)
    return tuple(repeat(x, 2))


def trunc_normal_(tensor, mean=0., std=1., a=-2., b=2.):
    return _tn(tensor, mean=mean, std=std, a=a, b=b)


def lecun_normal_(tensor):
    raise NotImplementedError("not on the ViPT hot path")


class DropPath(nn.Module):
    """eval-mode DropPath is the identity (timm 0.5.4 drop.py)."""
    def __init__(self, drop_prob=None):
        super().__init__()
        self.drop_prob = drop_prob

    def forward(self, x):
        assert not self.training
        return x


class Mlp(nn.Module):
    """timm 0.5.4 layers/mlp.py: fc1 -> act -> drop -> fc2 -> drop."""
    def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.):
        super().__init__()
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_features, out_features)
        self.drop = nn.Dropout(drop)

    def forward(self, x):
        x = self.fc1(x)
        x = self.act(x)
        x = self.drop(x)
        x = self.fc2(x)
        x = self.drop(x)
        return x
