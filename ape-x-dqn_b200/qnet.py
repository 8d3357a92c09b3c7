"""Stand-in Q-network used when the reference's duelling_network.py is not on sys.path
(e.g. on the GPU box).  Same architecture, parameter names and output triple as
duelling_network.py:3-28, so state dicts are interchangeable; the model itself is
outside the accelerated path and runs on cuDNN/cuBLAS.
"""
import torch
from torch import nn


def _conv_relu(c_in, c_out, kernel, stride):
    return nn.Sequential(nn.Conv2d(c_in, c_out, kernel, stride=stride), nn.ReLU())


def _fc_relu(n_in, n_out):
    return nn.Sequential(nn.Linear(n_in, n_out), nn.ReLU())


class DuellingQNet(nn.Module):
    FEATURES = 64 * 7 * 7

    def __init__(self, state_shape, action_dim, batch_coupled_mean=True):
        super().__init__()
        self.input_shape, self.action_dim = state_shape, action_dim
        # duelling_network.py:27 subtracts advantage.sum() over the WHOLE batch;
        # batch_coupled_mean=False gives the usual per-row mean instead.
        self.batch_coupled_mean = batch_coupled_mean
        specs = [(state_shape[0], 64, 8, 4), (64, 64, 4, 2), (64, 64, 3, 1)]
        for i, spec in enumerate(specs, start=1):
            setattr(self, "layer%d" % i, _conv_relu(*spec))
        self.value_stream_layer = _fc_relu(self.FEATURES, 512)
        self.advantage_stream_layer = _fc_relu(self.FEATURES, 512)
        self.value = nn.Linear(512, 1)
        self.advantage = nn.Linear(512, action_dim)

    def forward(self, x):
        for i in (1, 2, 3):
            x = getattr(self, "layer%d" % i)(x)
        feat = x.reshape(-1, self.FEATURES)
        v = self.value(self.value_stream_layer(feat))
        adv = self.advantage(self.advantage_stream_layer(feat))
        centre = adv.sum() if self.batch_coupled_mean else adv.sum(dim=1, keepdim=True)
        return v, adv, v + (adv - centre / self.action_dim)
